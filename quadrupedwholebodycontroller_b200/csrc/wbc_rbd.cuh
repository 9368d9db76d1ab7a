// Batched rigid-body dynamics of the 13-body ANYmal (base + 4 x {HIP,THIGH,SHANK}), one instance per warp.
//
// Replaces, for the 18-DoF floating-base ANYmal only, the ten iDynTree KinDynComputations calls of
// WholeBodyController::updateState (reference src/whole_body_controller.cpp:256-294):
//   setRobotState, getCenterOfMassPosition/Velocity, getFreeFloatingMassMatrix, getFrameFreeFloatingJacobian x4,
//   generalizedBiasForces - generalizedGravityForces, getWorldTransform x4, getFrameVel x4.
// Stage 1 ("leg phase"): lanes 0..3 each walk one leg (FK, CRBA, RNEA, Jacobian columns) in registers; the four
// legs only couple through the base, so each lane emits its leg's contribution expressed in the base frame.
// Stage 2 (wbc_assembly.cuh) combines them with all 32 lanes.
// Conventions are iDynTree's: 6-vectors [linear; angular], body-fixed link twists, MIXED representation outside.
#pragma once
#include "wbc_common.cuh"

namespace wbc {

// model constants (copied to shared memory once per warp); body b>=1 owns DOF b-1, leg l owns bodies 1+3l..3+3l
struct ModelConst {
    double restR[13*9], restP[13*3], axis[13*3], mass[13], mcom[13*3], inertia[13*9], footR[4*9], footP[4*3];
    double total_mass;
};
constexpr int MODEL_DOUBLES = sizeof(ModelConst)/sizeof(double);

// per-instance shared-memory scratch of the RBD + assembly stages
struct RbdScratch {
    // inputs
    double q[12], qd[12], pb[3], Rwb[9], twist[6], vbody[6];
    double pose_d[6], vel_d[6], acc_d[6], swp[12], swv[12], swa[12];
    double cmask[4];
    // leg-phase outputs (base frame)
    double crb[4][13];     // leg subtree composite inertia at the base origin: m, mc[3], I[9]
    double Fraw[12][6];    // raw (body-fixed) M[0:6, 6+j]
    double Hleg[4][9];     // joint-joint 3x3 block of leg l
    double plin[4][3];     // sum_k R_bk (I_k v_k).lin
    double fb[4][6];       // RNEA wrench the leg transmits to the base (gravity-free)
    double tauC[12];       // RNEA joint torques (gravity-free): C(q,nu) nu, joint part
    double footp[4][3];    // foot origin w.r.t. base origin, base coordinates
    double footv[4][3];    // foot origin linear velocity, base coordinates
    double jcol[12][3];    // d(foot lin vel)/d(qd_j), base coordinates
    // dense phase
    double mtot, comb[3], d[3], com[3], vcom[6], pose[6], Cb[6];
    double Mbb[36], Mbj[72], L[36], B[72], X[36], tmp66[36], Ttop[108], w6[6], vb6[6], W[6];
    double Jdv_st[12], Jdv_sw[12], rcmd[12], footpw[12], footvw[12], Tdot[108], Y[108];
};

// what ONE leg hands to the dense phase (67 doubles): the record wbc_rbd_legs_kernel writes per (robot, leg) and wbc_update_state_kernel reads back
struct LegOut { double crb[13], Fraw[3][6], Hleg[9], plin[3], fb[6], tauC[3], footp[3], footv[3], jcol[3][3]; };
constexpr int LEGOUT_DOUBLES = sizeof(LegOut)/sizeof(double);
// per-THREAD scratch of the leg kernel: same member names as RbdScratch, one leg wide (slot 0)
struct LegScratch {
    double q[3], qd[3], vbody[6];
    double crb[1][13], Fraw[3][6], Hleg[1][9], plin[1][3], fb[1][6], tauC[3], footp[1][3], footv[1][3], jcol[3][3];
};

// S = [0; a] (RevoluteJoint.cpp:189-196).  spatial inertia (m, mc, I about the origin) times motion vector
WBC_DEV void inertia_mul(double m,const double* mc,const double* I,const double* vl,const double* va,double* hl,double* ha){
    double t[3]; cross3(mc,va,t); hl[0]=m*vl[0]-t[0]; hl[1]=m*vl[1]-t[1]; hl[2]=m*vl[2]-t[2];
    double u[3],w[3]; cross3(mc,vl,u); mv3(I,va,w); ha[0]=u[0]+w[0]; ha[1]=u[1]+w[1]; ha[2]=u[2]+w[2];
}
// A_X_B applied to a force vector: lin' = R lin ; ang' = R ang + p x (R lin)
WBC_DEV void xf_force(const double* R,const double* p,double* fl,double* fa){
    double l[3],a[3],c[3]; mv3(R,fl,l); mv3(R,fa,a); cross3(p,l,c);
    fl[0]=l[0];fl[1]=l[1];fl[2]=l[2]; fa[0]=a[0]+c[0];fa[1]=a[1]+c[1];fa[2]=a[2]+c[2];
}
// (A_X_B)^-1 applied to a motion vector: ang' = R^T ang ; lin' = R^T (lin - p x ang)
WBC_DEV void xf_motion_inv(const double* R,const double* p,const double* vl,const double* va,double* ol,double* oa){
    double c[3],t[3]; cross3(p,va,c); t[0]=vl[0]-c[0];t[1]=vl[1]-c[1];t[2]=vl[2]-c[2]; mtv3(R,t,ol); mtv3(R,va,oa);
}
// inertia of B expressed in A for A_X_B = (R,p):  I_A = R I_B R^T - S(p)S(R mc) - S(R mc + m p)S(p)
WBC_DEV void xf_inertia(const double* R,const double* p,double m,const double* mc,const double* I,double* mcA,double* IA){
    double Rmc[3]; mv3(R,mc,Rmc);
    double RI[9],Rt[9]; for(int i=0;i<3;i++)for(int j=0;j<3;j++) Rt[3*i+j]=R[3*j+i];
    mm3(R,I,RI); mm3(RI,Rt,RI);
    mcA[0]=Rmc[0]+m*p[0]; mcA[1]=Rmc[1]+m*p[1]; mcA[2]=Rmc[2]+m*p[2];
    // S(a)S(b) = b a^T - (a.b) 1
    double pc=dot3(p,Rmc), ap=dot3(mcA,p);
    for(int i=0;i<3;i++)for(int j=0;j<3;j++){
        double t1=Rmc[i]*p[j]-(i==j?pc:0.0);      // S(p)S(Rmc)
        double t2=p[i]*mcA[j]-(i==j?ap:0.0);      // S(mcA)S(p)
        IA[3*i+j]=RI[3*i+j]-t1-t2;
    }
}
// Eigen::Quaterniond(w,x,y,z).toRotationMatrix() as used by floatingBaseStateCallback (wbc.cpp:209-213); no normalisation
WBC_DEV void quat_to_R(const double* q,double* R){
    double w=q[0],x=q[1],y=q[2],z=q[3];
    double tx=2*x,ty=2*y,tz=2*z, twx=tx*w,twy=ty*w,twz=tz*w, txx=tx*x,txy=ty*x,txz=tz*x, tyy=ty*y,tyz=tz*y,tzz=tz*z;
    R[0]=1-(tyy+tzz); R[1]=txy-twz; R[2]=txz+twy; R[3]=txy+twz; R[4]=1-(txx+tzz); R[5]=tyz-twx; R[6]=txz-twy; R[7]=tyz+twx; R[8]=1-(txx+tyy);
}
// Rodrigues (iDynTree Rotation::RotAxis)
WBC_DEV void rot_axis(const double* u,double th,double* r){
    double s,c; sincos(th,&s,&c); double t=1-c;
    r[0]=c+u[0]*u[0]*t;      r[1]=u[0]*u[1]*t-u[2]*s; r[2]=u[0]*u[2]*t+u[1]*s;
    r[3]=u[1]*u[0]*t+u[2]*s; r[4]=c+u[1]*u[1]*t;      r[5]=u[1]*u[2]*t-u[0]*s;
    r[6]=u[2]*u[0]*t-u[1]*s; r[7]=u[2]*u[1]*t+u[0]*s; r[8]=c+u[2]*u[2]*t;
}

// One leg: forward pass (ForwardKinematics.cpp:144-226), CRBA (Dynamics.cpp:146-273), RNEA dynamic phase
// (Dynamics.cpp:65-144) with zero gravity (== generalizedBiasForces - generalizedGravityForces, wbc.cpp:544-551),
// foot frame position / velocity / Jacobian columns (Jacobians.cpp:16-52).
// `slot` = index of the leg inside the scratch `s` (== leg for the per-robot RbdScratch, 0 for the per-thread LegScratch of wbc_rbd_legs_kernel).
template<class S> WBC_DEV void rbd_leg(const ModelConst& mc_,S& s,int leg,int slot){
    double Rpc[3][9], Rb[3][9], pbk[3][3], vl[3][3], va[3][3], al[3][3], aa[3][3];
    const double* vb=s.vbody;
    double pvl[3]={vb[0],vb[1],vb[2]}, pva[3]={vb[3],vb[4],vb[5]};
    double pal[3], paa[3]={0,0,0};
    { double c[3]; cross3(pva,pvl,c); pal[0]=-c[0]; pal[1]=-c[1]; pal[2]=-c[2]; }   // mixed zero acc -> body (KDC.cpp:240-261)
    double pR[9]={1,0,0,0,1,0,0,0,1}, pp[3]={0,0,0};
    for(int k=0;k<3;k++){
        const int b=1+3*leg+k; const double* ax=mc_.axis+3*b; const double* rp=mc_.restP+3*b;
        double rot[9]; rot_axis(ax,s.q[3*slot+k],rot);
        mm3(mc_.restR+9*b,rot,Rpc[k]);                                  // parent_X_child = rest * Rot(axis,q)
        mm3(pR,Rpc[k],Rb[k]);
        double t[3]; mv3(pR,rp,t); pbk[k][0]=t[0]+pp[0]; pbk[k][1]=t[1]+pp[1]; pbk[k][2]=t[2]+pp[2];
        xf_motion_inv(Rpc[k],rp,pvl,pva,vl[k],va[k]);
        const double qd=s.qd[3*slot+k];
        double vjA[3]={ax[0]*qd,ax[1]*qd,ax[2]*qd};
        va[k][0]+=vjA[0]; va[k][1]+=vjA[1]; va[k][2]+=vjA[2];
        xf_motion_inv(Rpc[k],rp,pal,paa,al[k],aa[k]);
        double c1[3],c2[3]; cross3(vl[k],vjA,c1); cross3(va[k],vjA,c2);   // v x (S qd)
        al[k][0]+=c1[0]; al[k][1]+=c1[1]; al[k][2]+=c1[2]; aa[k][0]+=c2[0]; aa[k][1]+=c2[1]; aa[k][2]+=c2[2];
        for(int i=0;i<9;i++) pR[i]=Rb[k][i];
        for(int i=0;i<3;i++){ pp[i]=pbk[k][i]; pvl[i]=vl[k][i]; pva[i]=va[k][i]; pal[i]=al[k][i]; paa[i]=aa[k][i]; }
    }
    // link momenta and RNEA link wrenches f = I a + v x* (I v)
    double fl[3][3], fa[3][3], pl[3]={0,0,0};
    for(int k=0;k<3;k++){
        const int b=1+3*leg+k;
        double hl[3],ha[3]; inertia_mul(mc_.mass[b],mc_.mcom+3*b,mc_.inertia+9*b,vl[k],va[k],hl,ha);
        double t[3]; mv3(Rb[k],hl,t); pl[0]+=t[0]; pl[1]+=t[1]; pl[2]+=t[2];
        inertia_mul(mc_.mass[b],mc_.mcom+3*b,mc_.inertia+9*b,al[k],aa[k],fl[k],fa[k]);
        double c1[3],c2[3],c3[3]; cross3(va[k],hl,c1); cross3(vl[k],hl,c2); cross3(va[k],ha,c3);
        for(int i=0;i<3;i++){ fl[k][i]+=c1[i]; fa[k][i]+=c2[i]+c3[i]; }
    }
    for(int i=0;i<3;i++) s.plin[slot][i]=pl[i];
    // backward: RNEA torques + composite inertias + mass-matrix entries
    double cm[3], cmc[3][3], cI[3][9];
    for(int k=2;k>=0;k--){
        const int b=1+3*leg+k;
        cm[k]=mc_.mass[b]; for(int i=0;i<3;i++) cmc[k][i]=mc_.mcom[3*b+i]; for(int i=0;i<9;i++) cI[k][i]=mc_.inertia[9*b+i];
        if(k<2){
            const int bc=b+1; double mcA[3],IA[9];
            xf_inertia(Rpc[k+1],mc_.restP+3*bc,cm[k+1],cmc[k+1],cI[k+1],mcA,IA);
            cm[k]+=cm[k+1]; for(int i=0;i<3;i++) cmc[k][i]+=mcA[i]; for(int i=0;i<9;i++) cI[k][i]+=IA[i];
            xf_force(Rpc[k+1],mc_.restP+3*bc,fl[k+1],fa[k+1]);
            for(int i=0;i<3;i++){ fl[k][i]+=fl[k+1][i]; fa[k][i]+=fa[k+1][i]; }
        }
        s.tauC[3*slot+k]=dot3(mc_.axis+3*b,fa[k]);
        // CRBA column of DOF (leg,k): F = Ic S
        double Fl[3],Fa[3],zero[3]={0,0,0}; inertia_mul(cm[k],cmc[k],cI[k],zero,mc_.axis+3*b,Fl,Fa);
        s.Hleg[slot][3*k+k]=dot3(mc_.axis+3*b,Fa);
        for(int a=k;a>=1;a--){
            xf_force(Rpc[a],mc_.restP+3*(1+3*leg+a),Fl,Fa);
            double h=dot3(mc_.axis+3*(3*leg+a),Fa);               // axis of body (leg,a-1)
            s.Hleg[slot][3*k+(a-1)]=h; s.Hleg[slot][3*(a-1)+k]=h;
        }
        xf_force(Rpc[0],mc_.restP+3*(1+3*leg),Fl,Fa);
        for(int i=0;i<3;i++){ s.Fraw[3*slot+k][i]=Fl[i]; s.Fraw[3*slot+k][3+i]=Fa[i]; }
    }
    { const int b=1+3*leg; double mcA[3],IA[9];
      xf_inertia(Rpc[0],mc_.restP+3*b,cm[0],cmc[0],cI[0],mcA,IA);
      s.crb[slot][0]=cm[0]; for(int i=0;i<3;i++) s.crb[slot][1+i]=mcA[i]; for(int i=0;i<9;i++) s.crb[slot][4+i]=IA[i];
      xf_force(Rpc[0],mc_.restP+3*b,fl[0],fa[0]);
      for(int i=0;i<3;i++){ s.fb[slot][i]=fl[0][i]; s.fb[slot][3+i]=fa[0][i]; } }
    // foot frame (fixed in SHANK): position, velocity of its origin, Jacobian columns, all in base coordinates
    double pf[3]; { double t[3]; mv3(Rb[2],mc_.footP+3*leg,t); pf[0]=t[0]+pbk[2][0]; pf[1]=t[1]+pbk[2][1]; pf[2]=t[2]+pbk[2][2]; }
    { double c[3],t[3]; cross3(va[2],mc_.footP+3*leg,c); t[0]=vl[2][0]+c[0]; t[1]=vl[2][1]+c[1]; t[2]=vl[2][2]+c[2];
      double w[3]; mv3(Rb[2],t,w); for(int i=0;i<3;i++){ s.footp[slot][i]=pf[i]; s.footv[slot][i]=w[i]; } }
    for(int k=0;k<3;k++){
        double ab[3],r[3]={pf[0]-pbk[k][0],pf[1]-pbk[k][1],pf[2]-pbk[k][2]},c[3];
        mv3(Rb[k],mc_.axis+3*(1+3*leg+k),ab); cross3(ab,r,c);
        for(int i=0;i<3;i++) s.jcol[3*slot+k][i]=c[i];
    }
}

}  // namespace wbc
