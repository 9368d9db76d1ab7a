// Centroidal transform + QP assembly, fused behind the RBD leg phase (same warp, same shared memory).
//
// Replaces WholeBodyController::updateState lines 260-293, computeTransformationMatrix, computeDerivatives,
// computeDesiredWrench, computeCommandedAccelerationSwingLegs, computeNonSlidingConstraints and the assembly half of
// solveQP (reference src/whole_body_controller.cpp:256-515).
// The reference forms T (18x18) and inverts it seven times per tick.  T = [[Xi, Xi B],[0, I]] with
// Xi = [[1, -S(d)],[0, 1]], d = p_com - p_base, B = Mbb^-1 Mbj, so T^-1 = [[X, -B],[0, I]] in closed form and
//   M_c = T^-T M T^-1 = blkdiag(X^T Mbb X, Mjj - Mjb B),   J_c = J T^-1 = [Jb X, Jj - Jb B],
//   bias_c,j = v_j - B^T v_b with v = C nu + M (Tdot^-1)_prev nu,   Tdot^-1 = -T^-1 Tdot T^-1 (top 6 rows only).
// Same numbers (to rounding), ~20x fewer flops, nothing leaves shared memory.
#pragma once
#include "wbc_rbd.cuh"

// The dense phase runs with WBC_ASM_NL lanes per robot (16: two robots share a warp).  Most of its loops have 6, 12 or 36 elements and its first block
// is serial, so a whole warp per robot kept 12 of 32 lanes busy (ncu, r02); `lane` below is the lane index INSIDE the robot's group.  Measured for
// 16384 robots (profiles/r02_update_state.txt): 32 lanes per robot 0.408 ms, 16 lanes 0.401 ms, 8 lanes 0.739 ms (57 KB of shared memory per warp leave
// 3 warps per SM: the kernel is latency bound and wants warps in flight more than it wants full lanes).
#if WBC_NL==1
  #define WBC_ASM_NL 1
  #define ASM_FOR(i,n) for(int i=lane;i<(n);i+=1)
#else
  #ifndef WBC_ASM_NL
  #define WBC_ASM_NL 16
  #endif
  #define ASM_FOR(i,n) _Pragma("unroll 1") for(int i=lane;i<(n);i+=WBC_ASM_NL)
#endif

namespace wbc {

// wbc.cpp:12-20
WBC_DEV void eul_rpy(const double* R,double* rpy){
    rpy[0]=atan2(R[7],R[8]); rpy[1]=atan2(-R[6],sqrt(R[7]*R[7]+R[8]*R[8])); rpy[2]=atan2(R[3],R[0]);
}

// Everything after the leg phase.  `st` is this instance's persistent state (global memory), `out` its compact QP.
// `first` != 0 resets the state like setInitialState (wbc.cpp:65-120): oldT = I, oldJ = 0, Tdot^-1 = 0, integral = 0.
// `plant` (nullable, SoA with stride N): what wbc_plant_step needs to turn the QP's centroidal accelerations back into generalized
// ones, nu_dot = T^-1 (a - Tdot nu):  d = p_com - p_base (3), B = Mbb^-1 Mbj (72), Tdot_top nu (6).
// `out` is the robot's QPData record (global memory in the kernel: every element of it is written exactly once and never read back here); `stg` holds the three blocks
// the assembly reads again after writing them -- Jc, Mbb, bias_j -- in shared memory (QPStage; the caller copies them to the record afterwards).  The host emulation
// passes the record itself for both.  (The whole record used to be staged: 5.7 KB of the 14.4 KB of shared memory per robot.)
struct QPStage { double Jc[12*18], Mbb[36], biasj[12]; };
template<class STG> WBC_FN void wbc_assemble(const ModelConst& mc_,const Params& P,RbdScratch& s,CtrlState& st,QPData& out,STG& stg,int lane,int switching,int first,double* plant,int N){
    const double* R=s.Rwb;
    // the previous tick's T (finite difference of D5) is requested now and used ~100 lines further down: inside the loop that consumes it every load
    // would sit behind the shared-memory store of the trip before (the compiler cannot prove the two do not alias) and cost a memory round trip per trip
    constexpr int KT=(108+WBC_ASM_NL-1)/WBC_ASM_NL; double told_[KT];
    #pragma unroll
    for(int k=0;k<KT;k++){ const int e=lane+WBC_ASM_NL*k; told_[k]=(e<108&&!first)?st.Told[e]:0.0; }
    constexpr int KI=(6+WBC_ASM_NL-1)/WBC_ASM_NL; double integ_[KI];
    #pragma unroll
    for(int k=0;k<KI;k++){ const int i=lane+WBC_ASM_NL*k; integ_[k]=(i<6&&!first)?st.integral[i]:0.0; }
    // ---- D1: base totals (locked inertia, CoM, momentum, gravity-free base wrench) ----
    if(lane==0){
        double m=mc_.mass[0], mcb[3]={mc_.mcom[0],mc_.mcom[1],mc_.mcom[2]}, I[9]; for(int i=0;i<9;i++) I[i]=mc_.inertia[i];
        for(int l=0;l<4;l++){ m+=s.crb[l][0]; for(int i=0;i<3;i++) mcb[i]+=s.crb[l][1+i]; for(int i=0;i<9;i++) I[i]+=s.crb[l][4+i]; }
        s.mtot=m; for(int i=0;i<3;i++) s.comb[i]=mcb[i]/m;
        double dw[3]; mv3(R,s.comb,dw); for(int i=0;i<3;i++){ s.d[i]=dw[i]; s.com[i]=dw[i]+s.pb[i]; }
        // raw locked inertia [[m1,-S(mc)],[S(mc),I]] -> MIXED: blkdiag(R,R) . blkdiag(R,R)^T   (KDC.cpp:2074-2152)
        double mcw[3]; mv3(R,mcb,mcw);
        double RI[9],Rt[9]; for(int i=0;i<3;i++)for(int j=0;j<3;j++) Rt[3*i+j]=R[3*j+i]; mm3(R,I,RI); mm3(RI,Rt,RI);
        const double S[9]={0,-mcw[2],mcw[1], mcw[2],0,-mcw[0], -mcw[1],mcw[0],0};
        for(int i=0;i<3;i++)for(int j=0;j<3;j++){ s.Mbb[6*i+j]=(i==j)?m:0.0; s.Mbb[6*i+3+j]=-S[3*i+j]; s.Mbb[6*(3+i)+j]=S[3*i+j]; s.Mbb[6*(3+i)+3+j]=RI[3*i+j]; }
        // CoM velocity = world-oriented total linear momentum / m  (KDC.cpp:1941-1958)
        double hl[3],ha[3]; inertia_mul(mc_.mass[0],mc_.mcom,mc_.inertia,s.vbody,s.vbody+3,hl,ha);
        double pl[3]={hl[0],hl[1],hl[2]}; for(int l=0;l<4;l++) for(int i=0;i<3;i++) pl[i]+=s.plin[l][i];
        double pw[3]; mv3(R,pl,pw); for(int i=0;i<3;i++){ s.vcom[i]=pw[i]/m; s.vcom[3+i]=s.twist[3+i]; }      // wbc.cpp:261
        double rpy[3]; eul_rpy(R,rpy); for(int i=0;i<3;i++){ s.pose[i]=s.com[i]; s.pose[3+i]=rpy[i]; }          // wbc.cpp:262-264
        // base row of C(q,nu)nu: f0 = I0 a0 + v0 x* (I0 v0) + sum legs, rotated to world (KDC.cpp:2653-2668)
        double a0l[3],a0a[3]={0,0,0}; { double c[3]; cross3(s.vbody+3,s.vbody,c); a0l[0]=-c[0];a0l[1]=-c[1];a0l[2]=-c[2]; }
        double fl[3],fa[3]; inertia_mul(mc_.mass[0],mc_.mcom,mc_.inertia,a0l,a0a,fl,fa);
        double c1[3],c2[3],c3[3]; cross3(s.vbody+3,hl,c1); cross3(s.vbody,hl,c2); cross3(s.vbody+3,ha,c3);
        for(int i=0;i<3;i++){ fl[i]+=c1[i]; fa[i]+=c2[i]+c3[i]; }
        for(int l=0;l<4;l++) for(int i=0;i<3;i++){ fl[i]+=s.fb[l][i]; fa[i]+=s.fb[l][3+i]; }
        double fw[3],tw[3]; mv3(R,fl,fw); mv3(R,fa,tw); for(int i=0;i<3;i++){ s.Cb[i]=fw[i]; s.Cb[3+i]=tw[i]; }
        // X = [[1,S(d)],[0,1]]
        const double Sd[9]={0,-s.d[2],s.d[1], s.d[2],0,-s.d[0], -s.d[1],s.d[0],0};
        for(int i=0;i<6;i++)for(int j=0;j<6;j++) s.X[6*i+j]=(i==j)?1.0:0.0;
        for(int i=0;i<3;i++)for(int j=0;j<3;j++) s.X[6*i+3+j]=Sd[3*i+j];
        // Cholesky Mbb = L L^T (6x6 SPD)
        for(int j=0;j<6;j++){
            double dd=s.Mbb[6*j+j]; for(int k=0;k<j;k++) dd-=s.L[6*j+k]*s.L[6*j+k];
            dd=sqrt(dd); s.L[6*j+j]=dd;
            for(int i=j+1;i<6;i++){ double v=s.Mbb[6*i+j]; for(int k=0;k<j;k++) v-=s.L[6*i+k]*s.L[6*j+k]; s.L[6*i+j]=v/dd; }
        }
    }
    WBC_SYNC();
    // ---- D2: Mbj (MIXED) column j = [R F.lin ; R F.ang]; foot world offsets/velocities; B = Mbb^-1 Mbj ----
    ASM_FOR(j,12){
        double a[3],b[3]; mv3(R,s.Fraw[j],a); mv3(R,s.Fraw[j]+3,b);
        double y[6]={a[0],a[1],a[2],b[0],b[1],b[2]};
        #pragma unroll
        for(int i=0;i<6;i++) s.Mbj[12*i+j]=y[i];
        #pragma unroll
        for(int i=0;i<6;i++){ double v=y[i];
            #pragma unroll
            for(int k=0;k<6;k++) if(k<i) v-=s.L[6*i+k]*y[k];
            y[i]=v/s.L[6*i+i]; }
        #pragma unroll
        for(int i=5;i>=0;i--){ double v=y[i];
            #pragma unroll
            for(int k=0;k<6;k++) if(k>i) v-=s.L[6*k+i]*y[k];
            y[i]=v/s.L[6*i+i]; }
        #pragma unroll
        for(int i=0;i<6;i++) s.B[12*i+j]=y[i];
    }
    ASM_FOR(l,4){
        double a[3],b[3]; mv3(R,s.footp[l],a); mv3(R,s.footv[l],b);
        for(int i=0;i<3;i++){ s.footpw[3*l+i]=a[i]+s.pb[i]; s.footvw[3*l+i]=b[i]; }
    }
    WBC_SYNC();
    // ---- D3: centroidal blocks ----
    // M_c,bb = X^T Mbb X
    ASM_FOR(e,36){ int i=e/6,j=e%6; double v=0; for(int k=0;k<6;k++) v+=s.Mbb[6*i+k]*s.X[6*k+j]; s.tmp66[e]=v; }
    WBC_SYNC();
    ASM_FOR(e,36){ int i=e/6,j=e%6; double v=0; for(int k=0;k<6;k++) v+=s.X[6*k+i]*s.tmp66[6*k+j]; stg.Mbb[e]=v; }
    // M_c,jj = Mjj - Mjb B   (Mjj is block diagonal per leg)
    ASM_FOR(e,144){ int i=e/12,j=e%12; double v=(i/3==j/3)?s.Hleg[i/3][3*(i%3)+(j%3)]:0.0;
        for(int k=0;k<6;k++) v-=s.Mbj[12*k+i]*s.B[12*k+j]; out.Mjj[e]=v; }
    // J_c = [Jb X, Jj - Jb B],  Jb(foot l) = [1, -S(r_l)], r_l = p_foot - p_base (world)   (Jacobians.cpp:16-52)
    ASM_FOR(e,216){ int i=e/18,j=e%18,l=i/3,r=i%3;
        double rl[3]={s.footpw[3*l]-s.pb[0],s.footpw[3*l+1]-s.pb[1],s.footpw[3*l+2]-s.pb[2]};
        // row r of [1, -S(r_l)] without a dynamically indexed local array
        const double jb[6]={ (r==0)?1.0:0.0, (r==1)?1.0:0.0, (r==2)?1.0:0.0,
                             (r==0)?-0.0:((r==1)?-rl[2]:rl[1]), (r==0)?rl[2]:((r==1)?-0.0:-rl[0]), (r==0)?-rl[1]:((r==1)?rl[0]:-0.0) };
        double v;
        if(j<6){ v=0; for(int k=0;k<6;k++) v+=jb[k]*s.X[6*k+j]; }
        else { int jj=j-6; v=0;
            if(jj/3==l){ const double* c=s.jcol[jj]; v=R[3*r]*c[0]+R[3*r+1]*c[1]+R[3*r+2]*c[2]; }
            for(int k=0;k<6;k++) v-=jb[k]*s.B[12*k+jj]; }
        stg.Jc[e]=v; }
    // T top rows = [Xi, Xi B], Xi = [[1,-S(d)],[0,1]]
    ASM_FOR(e,108){ int i=e/18,j=e%18; double v;
        if(j<6) v=(i==j)?1.0:((i<3&&j>=3)?-s.X[6*i+j]:0.0);
        else { int jj=j-6; v=s.B[12*i+jj]; if(i<3) for(int k=0;k<3;k++) v-=s.X[6*i+3+k]*s.B[12*(3+k)+jj]; }
        s.Ttop[e]=v; }
    // w6 = (Tdot^-1)_prev nu  (only its top 6 rows are non-zero)
    ASM_FOR(i,6){ double v=0; if(!first){ for(int k=0;k<6;k++) v+=st.TdotInv[18*i+k]*s.twist[k]; for(int k=0;k<12;k++) v+=st.TdotInv[18*i+6+k]*s.qd[k]; } s.w6[i]=v; }
    WBC_SYNC();
    // ---- D4: bias_c (joint part) = v_j - B^T v_b,  v = C nu + M w   (wbc.cpp:289) ----
    ASM_FOR(i,6){ double v=s.Cb[i]; for(int k=0;k<6;k++) v+=s.Mbb[6*i+k]*s.w6[k]; s.vb6[i]=v; }
    WBC_SYNC();
    ASM_FOR(j,12){ double v=s.tauC[j]; for(int k=0;k<6;k++) v+=s.Mbj[12*k+j]*s.w6[k]; for(int k=0;k<6;k++) v-=s.B[12*k+j]*s.vb6[k]; stg.biasj[j]=v; }
    // ---- D5: finite-difference terms -Jdot [v_c; qd] (wbc.cpp:384-402,504,507), then roll old <- current ----
    const double timeStep=1.0/P.loopRate;
    ASM_FOR(i,12){
        const double c=s.cmask[i/3]; double a=0,b=0;
        if(!switching){
            for(int k=0;k<18;k++){
                double jc=stg.Jc[18*i+k], vk=(k<6)?s.vcom[k]:s.qd[k-6];
                double ost=first?0.0:st.Jstold[18*i+k], osw=first?0.0:st.Jswold[18*i+k];
                a+=((c*jc-ost)/timeStep)*vk; b+=(((1.0-c)*jc-osw)/timeStep)*vk; }
        }
        s.Jdv_st[i]=a; s.Jdv_sw[i]=b;
    }
    WBC_SYNC();
    // Tdot^-1 (top rows) for the NEXT tick = -X [Tdot_b X, Tdot_j - Tdot_b B]   (wbc.cpp:293)
    #pragma unroll
    for(int k=0;k<KT;k++){ const int e=lane+WBC_ASM_NL*k; if(e<108){ int i=e/18,j=e%18; double v=0;
        if(!switching){ double told=first?((i==j)?1.0:0.0):told_[k]; v=(s.Ttop[e]-told)/timeStep; }
        s.Tdot[e]=v; } }
    WBC_SYNC();
    if(plant){
        ASM_FOR(i,3) plant[(size_t)i*N]=s.d[i];
        ASM_FOR(e,72) plant[(size_t)(3+e)*N]=s.B[e];
        ASM_FOR(i,6){ double v=0; for(int k=0;k<6;k++) v+=s.Tdot[18*i+k]*s.twist[k]; for(int k=0;k<12;k++) v+=s.Tdot[18*i+6+k]*s.qd[k]; plant[(size_t)(75+i)*N]=v; }
    }
    ASM_FOR(e,108){ int i=e/18,j=e%18; double v=0;
        if(j<6){ for(int k=0;k<6;k++) v+=s.Tdot[18*i+k]*s.X[6*k+j]; }
        else { v=s.Tdot[18*i+j]; for(int k=0;k<6;k++) v-=s.Tdot[18*i+k]*s.B[12*k+(j-6)]; }
        s.Y[e]=v; }
    WBC_SYNC();
    ASM_FOR(e,108){ int i=e/18,j=e%18; double v=0; for(int k=0;k<6;k++) v+=s.X[6*i+k]*s.Y[18*k+j]; st.TdotInv[e]=-v; st.Told[e]=s.Ttop[e]; }
    ASM_FOR(e,216){ const double c=s.cmask[e/54]; double jc=stg.Jc[e]; st.Jstold[e]=c*jc; st.Jswold[e]=(1.0-c)*jc; }
    // ---- D6: desired wrench (wbc.cpp:426-445), swing command (wbc.cpp:447-464) ----
    #pragma unroll
    for(int kk=0;kk<KI;kk++){ const int i=lane+WBC_ASM_NL*kk; if(i<6){
        const double kp=(i==2)?P.kpValueZ:P.kpValue;
        double mcb=0; for(int k=0;k<6;k++) mcb+=stg.Mbb[6*i+k]*s.acc_d[k];     // needs stg.Mbb: written before the D3 barrier
        const double gw=(i==2)?mc_.total_mass*GRAV:0.0;
        const double integ=integ_[kk];                                       // first tick: 0
        s.W[i]=-kp*(s.pose[i]-s.pose_d[i])-P.kdValue*(s.vcom[i]-s.vel_d[i])-P.kiValue*integ+gw+mcb;
        st.integral[i]=integ+(s.pose[i]-s.pose_d[i])/P.loopRate;
    } }
    ASM_FOR(i,12){
        double r=s.swa[i]+P.kdSwingValue*(s.swv[i]-s.footvw[i])+P.kpSwingValue*(s.swp[i]-s.footpw[i]);
        s.rcmd[i]=r*(1.0-s.cmask[i/3]);
    }
    WBC_SYNC();
    // ---- D7: QP vectors and the only dense H block (wbc.cpp:468-515) ----
    ASM_FOR(e,144){ int i=e/12,j=e%12; double ci=s.cmask[i/3],cj=s.cmask[j/3]; double v=0;
        for(int k=0;k<6;k++) v+=(ci*stg.Jc[18*i+k])*(cj*stg.Jc[18*j+k]);
        out.Hff[e]=v+((i==j)?1.0:0.0); }
    ASM_FOR(i,12){ double ci=s.cmask[i/3]; double v=0; for(int k=0;k<6;k++) v+=(ci*stg.Jc[18*i+k])*s.W[k]; out.gf[i]=-v; }
    const double INF=1.0e20;
    ASM_FOR(i,70){
        double lo,up;
        if(i<6){ const double gw=(i==2)?mc_.total_mass*GRAV:0.0; lo=-gw; up=-gw; }
        else if(i<18){ lo=-s.Jdv_st[i-6]; up=lo; }
        else if(i<34){ lo=-INF; up=0.0; }
        else if(i<46){ lo=-P.maxTorque-stg.biasj[i-34]; up=P.maxTorque-stg.biasj[i-34]; }
        else if(i<58){ lo=-INF; up=s.rcmd[i-46]-s.Jdv_sw[i-46]; }
        else { lo=s.rcmd[i-58]-s.Jdv_sw[i-58]; up=INF; }
        out.lbA[i]=lo; out.ubA[i]=up;
    }
    if(lane==0){ out.mu=P.friction; out.rho=P.slackWeight; for(int l=0;l<4;l++) out.cmask[l]=s.cmask[l]; }
    WBC_SYNC();
}

}  // namespace wbc
