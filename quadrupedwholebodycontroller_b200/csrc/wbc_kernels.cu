// CUDA kernels (sm_100a) + C ABI of the batched ANYmal whole-body controller.  See include/wbc_b200.h.
// One robot instance per warp, one warp per CTA (the solver keeps Q, the packed R/T triangle pair and all vectors of its
// instance in 41.5 KB of shared memory, 5 CTAs per SM), persistent CTAs pulling instances from an atomic work counter so
// that robots that need 100 working-set changes do not hold up robots that need none.  The closed loop around the
// controller (MotionPlanner tick, ideal-actuator plant step) runs as two more one-thread-per-robot kernels.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>
#include "../../include/wbc_b200.h"
#ifndef WBC_GANG                /* solver warps (= robots in flight) per CTA, one CTA per SM: as many as the 227 KB of shared memory hold */
  #if defined(WBC_FP32)
    #define WBC_GANG 11         /* 11 x 19.0 KB (vectors in shared memory) */
  #elif !defined(WBC_NO_TMEM)
    #define WBC_GANG 7          /* 7 x 32.4 KB, element-wise vectors in tensor memory */
  #else
    #define WBC_GANG 6          /* 6 x 37.6 KB */
  #endif
#endif
#define WBC_PARK_CHUNK 65536      /* robots per pair of launches of a parked cold batch: bounds the parking buffer at 2.7 GB */
#include "wbc_assembly.cuh"
#include "wbc_qp.cuh"
#include "wbc_planner.cuh"
#include "anymal_model.h"

using namespace wbc;

// ---------------------------------------------------------------- device-side batch views
struct BatchIn {      // SoA, instance index fastest
    const double *q,*qd,*pos,*quat,*twist,*pose,*vel,*acc,*swp,*swv,*swa;
    const unsigned char *contacts,*switching;
};
struct BatchOut { double *tau,*grf,*x,*com; int *rc,*nwsr; signed char* ws; };

__constant__ ModelConst c_model;

// ---------------------------------------------------------------- kernel 1a: rigid-body dynamics of the legs, one THREAD per (robot, leg)
// The four legs of a robot only couple through the base, so FK, CRBA, RNEA and the foot Jacobian columns of one leg are one independent
// register-resident job (wbc_rbd.cuh).  Run per warp-of-one-robot that job kept 4 of 32 lanes busy (r01: 10.2 active lanes on average); here
// a warp carries 8 robots x 4 legs = 32 such jobs.  Lane -> (robot, leg) with the leg fastest: the 3 q / qd loads of a warp touch 4 rows x 8
// consecutive robots of the SoA inputs (64-byte runs).  Output: one 67-double LegOut record per (robot, leg), read back coalesced by kernel 1b.
__global__ void __launch_bounds__(128) wbc_rbd_legs_kernel(int N,BatchIn in,LegOut* legs){
    __shared__ ModelConst mc;
    { const double* src=(const double*)&c_model; double* dst=(double*)&mc; for(int i=threadIdx.x;i<MODEL_DOUBLES;i+=blockDim.x) dst[i]=src[i]; }
    __syncthreads();
    const int t=blockIdx.x*blockDim.x+threadIdx.x, r=t>>2, leg=t&3;
    if(r>=N) return;
    LegScratch s;
    for(int k=0;k<3;k++){ s.q[k]=in.q[(size_t)(3*leg+k)*N+r]; s.qd[k]=in.qd[(size_t)(3*leg+k)*N+r]; }
    { double qu[4]={in.quat[r],in.quat[(size_t)N+r],in.quat[(size_t)2*N+r],in.quat[(size_t)3*N+r]}, R[9], tw[6];
      for(int k=0;k<6;k++) tw[k]=in.twist[(size_t)k*N+r];
      quat_to_R(qu,R); mtv3(R,tw,s.vbody); mtv3(R,tw+3,s.vbody+3); }                    // MIXED -> body-fixed (KDC.cpp:822-825), same operations as kernel 1b
    rbd_leg(mc,s,leg,0);
    LegOut& o=legs[(size_t)r*4+leg];
    for(int i=0;i<13;i++) o.crb[i]=s.crb[0][i];
    for(int k=0;k<3;k++) for(int i=0;i<6;i++) o.Fraw[k][i]=s.Fraw[k][i];
    for(int i=0;i<9;i++) o.Hleg[i]=s.Hleg[0][i];
    for(int i=0;i<3;i++){ o.plin[i]=s.plin[0][i]; o.tauC[i]=s.tauC[i]; o.footp[i]=s.footp[0][i]; o.footv[i]=s.footv[0][i]; }
    for(int i=0;i<6;i++) o.fb[i]=s.fb[0][i];
    for(int k=0;k<3;k++) for(int i=0;i<3;i++) o.jcol[k][i]=s.jcol[k][i];
}

// ---------------------------------------------------------------- kernel 1b: centroidal transform + QP assembly, 32 / WBC_ASM_NL = 2 robots per warp
// 16 lanes per robot (wbc_assembly.cuh); a warp takes consecutive robots, grid-stride.  Each group reads the four LegOut records of its robot
// (268 consecutive doubles), runs the dense phase in its own slice of shared memory and writes the compact QP.  A group past the end of the batch
// re-does the last robot in step with that robot's own group (same warp, same instructions, same values written): no divergent barriers.
struct UpdSmem { RbdScratch s; QPStage stg; };
constexpr int UPD_GROUPS=32/WBC_ASM_NL;
__global__ void __launch_bounds__(32) wbc_update_state_kernel(int N,Params prm,BatchIn in,const LegOut* legs,CtrlState* st,QPData* qpd,unsigned char* first,double* com,double* plantAux){
    extern __shared__ double smem[];
    ModelConst& mc=*reinterpret_cast<ModelConst*>(smem);
    const int lane=threadIdx.x&(WBC_ASM_NL-1), grp=threadIdx.x/WBC_ASM_NL;
    UpdSmem& u=reinterpret_cast<UpdSmem*>(reinterpret_cast<char*>(smem)+sizeof(ModelConst))[grp];
    RbdScratch& s=u.s; QPStage& stg=u.stg;
    { const double* src=(const double*)&c_model; double* dst=(double*)&mc; for(int i=threadIdx.x;i<MODEL_DOUBLES;i+=32) dst[i]=src[i]; }
    WBC_SYNC();
    for(int r0=blockIdx.x*UPD_GROUPS;r0<N;r0+=gridDim.x*UPD_GROUPS){
        const int r=(r0+grp<N)?r0+grp:N-1;
        // the controller state of this robot is read half-way through the assembly: ask L2 for it now
        if(lane==0) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(__cvta_generic_to_global(&st[r])),"r"((unsigned)sizeof(CtrlState)) : "memory");
        // every global load of the robot is issued before the first shared store (the scratch is reached through a generic pointer: a load behind a
        // store stays behind it, and each 16-element round would cost its own memory round trip)
        constexpr int K12=(12+WBC_ASM_NL-1)/WBC_ASM_NL, K6=(6+WBC_ASM_NL-1)/WBC_ASM_NL, KL=(4*LEGOUT_DOUBLES+WBC_ASM_NL-1)/WBC_ASM_NL;
        double v12[5][K12], v6[4][K6], vpb=0.0, q0=1.0, q1=0.0, q2=0.0, q3=0.0, tl[KL]; unsigned char vc=0;
        #pragma unroll
        for(int j=0;j<K12;j++){ const int k=lane+WBC_ASM_NL*j; const bool in_=k<12; const size_t o=(size_t)(in_?k:0)*N+r;
            v12[0][j]=in_?in.q[o]:0.0; v12[1][j]=in_?in.qd[o]:0.0; v12[2][j]=in_?in.swp[o]:0.0; v12[3][j]=in_?in.swv[o]:0.0; v12[4][j]=in_?in.swa[o]:0.0; }
        #pragma unroll
        for(int j=0;j<K6;j++){ const int k=lane+WBC_ASM_NL*j; const bool in_=k<6; const size_t o=(size_t)(in_?k:0)*N+r;
            v6[0][j]=in_?in.twist[o]:0.0; v6[1][j]=in_?in.pose[o]:0.0; v6[2][j]=in_?in.vel[o]:0.0; v6[3][j]=in_?in.acc[o]:0.0; }
        if(lane<3) vpb=in.pos[(size_t)lane*N+r];
        if(lane<4) vc=in.contacts[(size_t)lane*N+r];
        if(lane==0){ q0=in.quat[r]; q1=in.quat[N+r]; q2=in.quat[2*N+r]; q3=in.quat[3*N+r]; }
        { const double* src=(const double*)&legs[(size_t)r*4];                            // leg-phase results of kernel 1a
          #pragma unroll
          for(int j=0;j<KL;j++){ const int i=lane+WBC_ASM_NL*j; tl[j]=(i<4*LEGOUT_DOUBLES)?src[i]:0.0; } }
        #pragma unroll
        for(int j=0;j<K12;j++){ const int k=lane+WBC_ASM_NL*j; if(k<12){ s.q[k]=v12[0][j]; s.qd[k]=v12[1][j]; s.swp[k]=v12[2][j]; s.swv[k]=v12[3][j]; s.swa[k]=v12[4][j]; } }
        #pragma unroll
        for(int j=0;j<K6;j++){ const int k=lane+WBC_ASM_NL*j; if(k<6){ s.twist[k]=v6[0][j]; s.pose_d[k]=v6[1][j]; s.vel_d[k]=v6[2][j]; s.acc_d[k]=v6[3][j]; } }
        if(lane<3) s.pb[lane]=vpb;
        if(lane<4) s.cmask[lane]=vc?1.0:0.0;
        if(lane==0){ double qu[4]={q0,q1,q2,q3}; quat_to_R(qu,s.Rwb); }
        #pragma unroll
        for(int j=0;j<KL;j++){ const int i=lane+WBC_ASM_NL*j; if(i<4*LEGOUT_DOUBLES){ const int l=i/LEGOUT_DOUBLES, o=i-l*LEGOUT_DOUBLES; const double v=tl[j];       // -> the scratch fields the dense phase reads
              if(o<13) s.crb[l][o]=v; else if(o<31) (&s.Fraw[3*l][0])[o-13]=v; else if(o<40) s.Hleg[l][o-31]=v; else if(o<43) s.plin[l][o-40]=v;
              else if(o<49) s.fb[l][o-43]=v; else if(o<52) s.tauC[3*l+o-49]=v; else if(o<55) s.footp[l][o-52]=v; else if(o<58) s.footv[l][o-55]=v;
              else (&s.jcol[3*l][0])[o-58]=v; } }
        WBC_SYNC();
        if(lane==0){ mtv3(s.Rwb,s.twist,s.vbody); mtv3(s.Rwb,s.twist+3,s.vbody+3); }   // MIXED -> body-fixed (KDC.cpp:822-825)
        WBC_SYNC();
        wbc_assemble(mc,prm,s,st[r],qpd[r],stg,lane,(int)in.switching[r],(int)first[r],plantAux?plantAux+r:nullptr,N);
        WBC_SYNC();
        { QPData& q=qpd[r]; ASM_FOR(i,216) q.Jc[i]=stg.Jc[i]; ASM_FOR(i,36) q.Mbb[i]=stg.Mbb[i]; ASM_FOR(i,12) q.biasj[i]=stg.biasj[i]; }
        ASM_FOR(k,3) com[k*N+r]=s.com[k];
        WBC_SYNC();
    }
}

// ---------------------------------------------------------------- kernel 2a: TQ rebuild of the warm start, ahead of the solve
// The first half of SQProblem::setupNewAuxiliaryQP -- Q = I, then ~20 Givens re-adds of the previous active constraints -- is most of a quiet
// control tick, needs only Q, T, the index lists and four 42-vectors (QPWORK_TQ_BYTES = 23.8 KB: 9 robots per SM instead of the 5 the whole
// workspace allows) and is 1.5 k instructions of code.  It runs here for every robot that has a previous solution and leaves Q, T and the
// lists in HBM (QPPre, 22 KB per robot); wbc_solve_qp_kernel picks them up where the single-kernel path would have built them.  Same
// functions, same arithmetic, same order.  One robot per CTA (the hardware queue balances: every robot costs about the same here).
__global__ void __launch_bounds__(32) wbc_warm_setup_kernel(int N,const QPData* qpd,const QPState* qs,const unsigned char* first,QPPre* pre,int canary){
    extern __shared__ double smem[];
    QPWork& w=*reinterpret_cast<QPWork*>(smem);                       // ONLY the first QPWORK_TQ_BYTES exist
    const int lane=threadIdx.x, r=blockIdx.x;
    if(r>=N||first[r]!=0) return;
    // self-check (WBC_QP_SETUP_CANARY=1, tests): launched with the WHOLE workspace, everything past the prefix is filled with a NaN pattern first and
    // must still hold it at the end; a read from there poisons the result, a write is reported by poisoning it on purpose (see the end)
    unsigned long long* guard=reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(smem)+QPWORK_TQ_BYTES);
    constexpr int NGUARD=(int)((sizeof(QPWork)-QPWORK_TQ_BYTES)/8);
    if(canary){ LANE_FOR(i,NGUARD) guard[i]=0xFFFFFFFFFFFFFFFFull; WBC_SYNC(); }
    QPCtx c; c.lane=lane; c.nFR=c.nFX=c.nAC=c.nIAC=0; c.frIdentity=0; c.infeasible=c.unbounded=0; c.tau=0.0; c.rampOffset=0; c.status=QPS_NOTINITIALISED; c.qg=&qpd[r]; c.tm=0u; c.tqBar=0u; c.tqParity=0u;
#if defined(WBC_PROFILE)
    c.flops=0.0;
#endif
    const QPState& S=qs[r];
    signed char* old=reinterpret_cast<signed char*>(w.tmpv);
    { signed char t[3];
      #pragma unroll
      for(int k=0;k<3;k++){ const int i=lane+32*k; t[k]=(i<70)?S.cstat[i]:(signed char)0; }
      #pragma unroll
      for(int k=0;k<3;k++){ const int i=lane+32*k; if(i<70) old[i]=t[k]; } }
    WBC_SYNC();
    rebuild_TQ(w,c,qpd[r],old);
    WBC_SYNC();
    store_prebuilt_TQ(w,c,pre[r]);
    if(canary){ bool bad=false; LANE_FOR(i,NGUARD) if(guard[i]!=0xFFFFFFFFFFFFFFFFull) bad=true;
                if(wbc_ballot(bad)!=0u){ WBC_SYNC(); LANE_FOR(i,42*LD) pre[r].Q[i]=__longlong_as_double(-1ll); } }
}

// ---------------------------------------------------------------- kernel 2: active-set QP solve (init / hotstart) + torque recovery
// One robot per warp; WBC_GANG warps per CTA (one CTA per SM: 7 x 32.4 KB of shared memory + 7 x 7.3 KB of tensor memory in fp64, 11 x 19.0 KB in fp32) walk the iterations of their robots in
// step (wbc_qp.cuh "gang mode").  Robots are pulled from an atomic work counter.  cohort = 0: a warp fetches its next robot as soon as
// it is done (hot-started ticks: iteration counts 0..100); cohort = 1: the gang fetches together and waits for its slowest member (cold
// batches: all members then sit at the same iteration index with matrices of the same size, which keeps the barrier waits short).
#define WBC_QP_THREADS (32*WBC_GANG)
#if WBC_LV_TMEM
constexpr int WBC_TM_NEED_COLS=((WBC_GANG+3)/4)*LV_SLOTS*(int)WBC_TM_COLS_PER_REAL;      // column blocks of the robots that share a lane quarter
constexpr int WBC_TM_ALLOC_COLS= WBC_TM_NEED_COLS<=32?32 : WBC_TM_NEED_COLS<=64?64 : WBC_TM_NEED_COLS<=128?128 : WBC_TM_NEED_COLS<=256?256 : 512;
static_assert(WBC_TM_NEED_COLS<=512&&WBC_GANG<=16,"tensor memory: 512 columns, 4 lane quarters");
static_assert((size_t)WBC_GANG*(sizeof(QPWork)+16)+16<=232448,"gang workspace exceeds the 227 KB of shared memory a CTA can have");
#endif
__global__ void __launch_bounds__(WBC_QP_THREADS,1) wbc_solve_qp_kernel(int N,const QPData* qpd,QPState* qs,unsigned char* first,BatchOut out,int* counter,int nWSRmax,long long* prof,int cohort,const QPPre* pre,int* stats){
    extern __shared__ double smem[];
    QPWork& w=*reinterpret_cast<QPWork*>(reinterpret_cast<char*>(smem)+(size_t)(threadIdx.x>>5)*sizeof(QPWork));
    const int lane=threadIdx.x&31;
    // two mbarriers per warp behind the workspaces: [0] the QP blocks (fp64 build), [1] the rebuilt Q / T of a hot-started robot
    const unsigned barH=wbc_saddr(reinterpret_cast<char*>(smem)+(size_t)WBC_GANG*sizeof(QPWork)+16*(threadIdx.x>>5)), barQ=barH+8u;
    unsigned parH=0u, parQ=0u;
    if(lane==0){ wbc_mbar_init(barH,1u); wbc_mbar_init(barQ,1u); wbc_mbar_init_fence(); wbc_fence_proxy_async(); }
    WBC_SYNC();
#if WBC_LV_TMEM
    // tensor memory for the lane-owned vectors of the gang's robots: warp w uses lane quarter w % 4 (the only one its tcgen05.ld / st can reach) and the
    // column block w / 4.  One CTA per SM: the allocation cannot contend.
    unsigned* tmBaseS=reinterpret_cast<unsigned*>(reinterpret_cast<char*>(smem)+(size_t)WBC_GANG*sizeof(QPWork)+16*WBC_GANG);
    if((threadIdx.x>>5)==0){
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(wbc_saddr(tmBaseS)),"n"(WBC_TM_ALLOC_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmRobot=*tmBaseS+((((unsigned)(threadIdx.x>>5)&3u)*32u)<<16)+(unsigned)(threadIdx.x>>7)*(unsigned)(LV_SLOTS*WBC_TM_COLS_PER_REAL);
#endif
    for(;;){
        int r=0; if(lane==0) r=atomicAdd(counter,1); r=__shfl_sync(0xffffffffu,r,0);
        if(cohort){ if(gang_bar(r>=N)) break; }                      // the whole gang is out of work
        if(r>=N){ while(!gang_bar(1)){} if(cohort) continue; break; }     // keep the others' barriers complete until every warp is done
        const bool isFirst=first[r]!=0;
        QPCtx c; c.lane=lane; c.nFR=c.nFX=c.nAC=c.nIAC=0; c.frIdentity=0; c.infeasible=c.unbounded=0; c.tau=0.0; c.rampOffset=0; c.status=QPS_NOTINITIALISED; c.qg=&qpd[r];
        c.tqBar=0u; c.tqParity=0u;
#if WBC_LV_TMEM
        c.tm=tmRobot;
#else
        c.tm=0u;
#endif
        // Bulk asynchronous copies first (cp.async.bulk -> mbarrier): the robot's QP blocks and, for a hot-started robot whose factorisation kernel 2a
        // rebuilt, Q and T (21.7 KB in fp64) -- the latter is not needed before qp_hotstart has formed A x and the auxiliary gradient, so its
        // flight time is hidden.  WBC_SYNC + proxy fence: every lane has finished with the previous robot's workspace before the async proxy writes it.
        WBC_SYNC();
        const bool bulkQT=(pre!=nullptr)&&!isFirst;
        if(lane==0){
            wbc_fence_proxy_async();
#if !defined(WBC_FP32)
            wbc_mbar_expect_tx(barH,(unsigned)(QPHOT_HEAD*sizeof(double)));
            wbc_bulk_g2s(wbc_saddr(&w.qp),&qpd[r],(unsigned)(QPHOT_HEAD*sizeof(double)),barH);
#endif
            if(bulkQT){ constexpr unsigned QTB=(unsigned)(((42*LD+TRI42+1)*sizeof(real))/16*16);
                wbc_mbar_expect_tx(barQ,QTB); wbc_bulk_g2s(wbc_saddr(w.Q),pre[r].Q,QTB,barQ); }
        }
        if(bulkQT){ c.tqBar=barQ; c.tqParity=parQ; parQ^=1u; }
        // all global loads of a robot are issued before the first shared store: the workspace is reached through a generic pointer, so the
        // compiler will not move a load above a store by itself and every 32-element round would cost a full memory round trip.
        // Only the matrix blocks (head of QPData) and mu, rho, cmask (its last 6 elements) go to shared memory, in the solver's type.
#if defined(WBC_FP32)
        { const double* src=(const double*)&qpd[r]; real* dst=(real*)&w.qp; constexpr int K=(QPHOT_HEAD+6+31)/32; double t[K];
          #pragma unroll
          for(int k=0;k<K;k++){ const int i=lane+32*k; t[k]=(i<QPHOT_HEAD)?__ldg(src+i):((i<QPHOT_HEAD+6)?__ldg(src+(QPDATA_DOUBLES-6)+(i-QPHOT_HEAD)):0.0); }
          #pragma unroll
          for(int k=0;k<K;k++){ const int i=lane+32*k; if(i<QPHOT_HEAD+6) dst[i]=(real)t[k]; } }
#else
        const double tailv=(lane<6)?__ldg((const double*)&qpd[r]+(QPDATA_DOUBLES-6)+lane):0.0;
#endif
#if defined(WBC_PROFILE)
        for(int k=0;k<8;k++) c.prof[k]=0;
        c.flops=0.0;
        const long long tstart=clock64();
#endif
        QPState& S=qs[r];
        if(!isFirst){
            real tx[2],tl[2],tu[2],ty[4],ta[3],tb[3]; signed char sb[2],sc[3];
            #pragma unroll
            for(int k=0;k<2;k++){ const int i=lane+32*k; const bool in=i<42; tx[k]=in?S.x[i]:real(0); tl[k]=in?S.lb[i]:real(0); tu[k]=in?S.ub[i]:real(0); sb[k]=in?S.bstat[i]:(signed char)0; }
            #pragma unroll
            for(int k=0;k<4;k++){ const int i=lane+32*k; ty[k]=(i<112)?S.y[i]:real(0); }
            #pragma unroll
            for(int k=0;k<3;k++){ const int i=lane+32*k; const bool in=i<70; ta[k]=in?S.Ax_l[i]:real(0); tb[k]=in?S.Ax_u[i]:real(0); sc[k]=in?S.cstat[i]:(signed char)0; }
            c.rampOffset=S.rampOffset; c.status=S.status;
            LV_HOIST();
            #pragma unroll
            for(int k=0;k<2;k++){ const int i=lane+32*k; if(i<42){ w.x[i]=tx[k]; w.bstat[i]=sb[k]; } if(WBC_LV_TMEM||i<42){ LV_ST(lb,i,k,tl[k]); LV_ST(ub,i,k,tu[k]); } }
            #pragma unroll
            for(int k=0;k<4;k++){ const int i=lane+32*k; if(i<112) w.y[i]=ty[k]; }
            #pragma unroll
            for(int k=0;k<3;k++){ const int i=lane+32*k; if(i<70) w.cstat[i]=sc[k]; if(WBC_LV_TMEM||i<70){ LV_ST(Ax_l,i,k,ta[k]); LV_ST(Ax_u,i,k,tb[k]); } }
            LV_DONE();
        }
#if !defined(WBC_FP32)
        if(lane<6) ((real*)&w.qp)[QPHOT_HEAD+lane]=tailv;
        wbc_mbar_wait(barH,parH); parH^=1u;
#endif
        WBC_SYNC();
        int nWSR=nWSRmax;
        const int rcv= isFirst? qp_init<false>(w,c,nWSR) : qp_hotstart(w,c,nWSR,pre?&pre[r]:nullptr);
        if(c.tqBar){ wbc_mbar_wait(c.tqBar,c.tqParity); c.tqBar=0u; }            // qp_hotstart returned before it needed Q / T: the copy must still land before the workspace is reused
        WBC_SYNC();
        // QProblemB::getPrimalSolution (QProblemB.cpp:774-793) hands x out only in these three states; otherwise the caller's qpSolution_ keeps
        // its previous value (wbc.cpp:536) -- and so does out.x.  The solver's own iterate S.x is stored regardless (it IS the SQProblem member).
        const bool haveSol=(c.status==QPS_AUXILIARYQPSOLVED||c.status==QPS_HOMOTOPYQPSOLVED||c.status==QPS_SOLVED);
        LV_FOR(i,sl,42){ LV_LD(lbi,lb,i,sl); LV_LD(ubi,ub,i,sl); LV_WAIT(); LV_USE(lbi); LV_USE(ubi);
            if(LV_IN(i,42)){ S.x[i]=w.x[i]; S.lb[i]=lbi; S.ub[i]=ubi; S.bstat[i]=w.bstat[i]; if(haveSol) out.x[i*N+r]=w.x[i]; else w.x[i]=(real)out.x[i*N+r]; out.ws[i*N+r]=w.bstat[i]; } }
        LANE_FOR(i,112) S.y[i]=w.y[i];
        LV_FOR(i,sl,70){ LV_LD(al,Ax_l,i,sl); LV_LD(au,Ax_u,i,sl); LV_WAIT(); LV_USE(al); LV_USE(au);
            if(LV_IN(i,70)){ S.Ax_l[i]=al; S.Ax_u[i]=au; S.cstat[i]=w.cstat[i]; out.ws[(42+i)*N+r]=w.cstat[i]; } }
        WBC_SYNC();
        // computeJointTorques (wbc.cpp:553-577) on the solution still in shared memory: tau = M_c,jj qdd* + bias_c,j - Jst_c,j^T f*, f* = x[18:30];
        // a robot whose solve did not return 0 publishes zero torques (terminate(), :626-635).  w.x = qpSolution_ here (see above).
        if(lane<12){ const int i=lane; const QPHotR& q=WQP;
            real s=0; for(int j=0;j<12;j++) s+=q.Mjj[12*i+j]*w.x[6+j];
            real t=0; for(int j=0;j<12;j++) t+=(q.cmask[j/3]*q.Jc[18*j+6+i])*w.x[18+j];
            out.tau[i*N+r]= (rcv==0)? (double)(s+(real)c.qg->biasj[i]-t) : 0.0;
            out.grf[i*N+r]= w.x[18+i]; }
        if(lane==0){ S.rampOffset=c.rampOffset; S.status=c.status; out.rc[r]=rcv; out.nwsr[r]=nWSR; first[r]=0;
                     if(stats){ stats[r]+=1; if(rcv!=0) stats[N+r]+=1; stats[2*N+r]+=nWSR; if(nWSR>stats[3*N+r]) stats[3*N+r]=nWSR; }   // ticks, failed ticks, sum / max of nWSR
#if defined(WBC_PROFILE)
                     if(prof){ c.prof[6]=clock64()-tstart; for(int k=0;k<8;k++) prof[10*(size_t)r+k]=c.prof[k]; prof[10*(size_t)r+8]=(long long)c.flops; prof[10*(size_t)r+9]=isFirst?1:0; }
#endif
        }
        WBC_SYNC();
        if(cohort){ while(!gang_bar(1)){} }                           // drain until every member of the gang is done
    }
#if WBC_LV_TMEM
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();                                                  // every warp left the loop at the same barrier: nobody uses tensor memory any more
    if((threadIdx.x>>5)==0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(*tmBaseS),"n"(WBC_TM_ALLOC_COLS) : "memory");
#endif
}

// ---------------------------------------------------------------- kernel 2c: all-cold batches, with PARKING
// When every robot of the batch cold-starts, kernel 2 fetches gang by gang (all members then sit at the same iteration index, which keeps the barrier waits short) -- and
// every gang idles until its slowest member is done: iteration counts 62..100 inside a gang of 7 cost 11 % of the barrier rounds of cold4096.  This kernel runs such a
// batch as two launches over the robots [rBase, rEnd):
//   PHASE 1: gang fetch; as soon as the first member of a gang has finished (and the cheap early iterations are over, PARK_MIN_ITER) the others stop at the top of their
//            next iteration and PARK: workspace, lane-owned vectors and context go to park[k] (k from parkCtl[0]), the robot's number to parkList[k];
//   PHASE 2: per-warp fetch over the parked solves (counter = parkCtl[1]); each is restored and continued from exactly where it stopped.
// Results are bit-identical to kernel 2.  It is a kernel of its own (cold starts only: no hot-start code below it) so that kernel 2 and the functions ptxas compiles
// under it stay exactly what they were: register allocation runs over a kernel's whole call tree, and growing kernel 2's body by the parking code made 25 of its 45
// device functions longer and every launch 3 % slower (profiles/r02_gang_variants.txt, r02p).
WBC_DEV bool gang_all(int done){ return gang_count(done)==(int)blockDim.x; }
__device__ __noinline__ void park_store(const QPWork& w,const QPCtx& c,QPPark* park,int* parkCtl,int* parkList,int r,int lane){
    int slot=0; if(lane==0) slot=atomicAdd(&parkCtl[0],1); slot=__shfl_sync(0xffffffffu,slot,0);
    QPPark& P=park[slot];
    WBC_SYNC();
    { const double* src=reinterpret_cast<const double*>(&w); double* dst=reinterpret_cast<double*>(&P.w);
      for(int i=lane;i<(int)(sizeof(QPWork)/sizeof(double));i+=32) dst[i]=src[i]; }
#if WBC_LV_TMEM
    { real v[LV_SLOTS];
      #pragma unroll
      for(int s2=0;s2<LV_SLOTS;s2++) wbc_tm_ld(c.tm+WBC_TM_COLS_PER_REAL*(unsigned)s2,v[s2]);
      wbc_tm_wait_ld();
      #pragma unroll
      for(int s2=0;s2<LV_SLOTS;s2++){ wbc_tm_use(v[s2]); P.lv[s2*32+lane]=v[s2]; } }
#endif
    if(lane==0){ P.c=c; parkList[slot]=r; }
    WBC_SYNC();
}
// the workspace by one bulk copy (issued here, completed on `bar` by the caller), the lane-owned vectors into tensor memory, the context into every lane
__device__ __noinline__ void park_restore(QPWork& w,QPCtx& c,const QPPark& P,unsigned bar,int lane,const QPData* qg,unsigned tm){
    if(lane==0){ wbc_fence_proxy_async(); wbc_mbar_expect_tx(bar,(unsigned)sizeof(QPWork)); wbc_bulk_g2s(wbc_saddr(&w),&P.w,(unsigned)sizeof(QPWork),bar); }
    c=P.c; c.lane=lane; c.qg=qg; c.tqBar=0u; c.tqParity=0u; c.parkOK=0; c.resuming=1; c.tm=tm;
#if WBC_LV_TMEM
    { real v[LV_SLOTS];
      #pragma unroll
      for(int s2=0;s2<LV_SLOTS;s2++) v[s2]=P.lv[s2*32+lane];
      #pragma unroll
      for(int s2=0;s2<LV_SLOTS;s2++) wbc_tm_st(tm+WBC_TM_COLS_PER_REAL*(unsigned)s2,v[s2]);
      wbc_tm_wait_st(); }
#endif
}
template<int PHASE> __global__ void __launch_bounds__(WBC_QP_THREADS,1) wbc_solve_cold_kernel(int N,const QPData* qpd,QPState* qs,unsigned char* first,BatchOut out,int* counter,int nWSRmax,long long* prof,int* stats,
                                                                                          int rBase,int rEnd,QPPark* park,int* parkCtl,int* parkList){
    extern __shared__ double smem[];
    QPWork& w=*reinterpret_cast<QPWork*>(reinterpret_cast<char*>(smem)+(size_t)(threadIdx.x>>5)*sizeof(QPWork));
    const int lane=threadIdx.x&31;
    const unsigned barH=wbc_saddr(reinterpret_cast<char*>(smem)+(size_t)WBC_GANG*sizeof(QPWork)+16*(threadIdx.x>>5));
    unsigned parH=0u;
    if(lane==0){ wbc_mbar_init(barH,1u); wbc_mbar_init(barH+8u,1u); wbc_mbar_init_fence(); wbc_fence_proxy_async(); }
    WBC_SYNC();
    unsigned tmRobot=0u;
#if WBC_LV_TMEM
    unsigned* tmBaseS=reinterpret_cast<unsigned*>(reinterpret_cast<char*>(smem)+(size_t)WBC_GANG*sizeof(QPWork)+16*WBC_GANG);
    if((threadIdx.x>>5)==0){
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(wbc_saddr(tmBaseS)),"n"(WBC_TM_ALLOC_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    tmRobot=*tmBaseS+((((unsigned)(threadIdx.x>>5)&3u)*32u)<<16)+(unsigned)(threadIdx.x>>7)*(unsigned)(LV_SLOTS*WBC_TM_COLS_PER_REAL);
#endif
    constexpr bool cohort=(PHASE==1);
    for(;;){
        int r=0; if(lane==0) r=atomicAdd(counter,1); r=__shfl_sync(0xffffffffu,r,0);
        int pk=-1;
        if(PHASE==2){ pk=r; r= pk<parkCtl[0] ? parkList[pk] : N; }
        else { r+=rBase; if(r>=rEnd) r=N; }
        if(cohort){ if(gang_all(r>=N)) break; }
        if(r>=N){ while(!gang_all(1)){} if(cohort) continue; break; }
        QPCtx c; c.lane=lane; c.nFR=c.nFX=c.nAC=c.nIAC=0; c.frIdentity=0; c.infeasible=c.unbounded=0; c.tau=0.0; c.rampOffset=0; c.status=QPS_NOTINITIALISED; c.qg=&qpd[r];
        c.tqBar=0u; c.tqParity=0u; c.tm=tmRobot; c.parkOK=(PHASE==1)?1:0; c.resuming=0; c.park_iter=0; c.park_nWSRmax=0; c.park_farbound=real(0.0);
        WBC_SYNC();                                                   // every lane has finished with the previous robot's workspace
        if(PHASE==2){
            park_restore(w,c,park[pk],barH,lane,&qpd[r],tmRobot);
            wbc_mbar_wait(barH,parH); parH^=1u;
        } else {
            // the robot's QP blocks: bulk copy (fp64) / conversion (fp32), like kernel 2
#if defined(WBC_FP32)
            { const double* src=(const double*)&qpd[r]; real* dst=(real*)&w.qp; constexpr int K=(QPHOT_HEAD+6+31)/32; double t[K];
              #pragma unroll
              for(int k=0;k<K;k++){ const int i=lane+32*k; t[k]=(i<QPHOT_HEAD)?__ldg(src+i):((i<QPHOT_HEAD+6)?__ldg(src+(QPDATA_DOUBLES-6)+(i-QPHOT_HEAD)):0.0); }
              #pragma unroll
              for(int k=0;k<K;k++){ const int i=lane+32*k; if(i<QPHOT_HEAD+6) dst[i]=(real)t[k]; } }
#else
            if(lane==0){ wbc_fence_proxy_async(); wbc_mbar_expect_tx(barH,(unsigned)(QPHOT_HEAD*sizeof(double))); wbc_bulk_g2s(wbc_saddr(&w.qp),&qpd[r],(unsigned)(QPHOT_HEAD*sizeof(double)),barH); }
            const double tailv=(lane<6)?__ldg((const double*)&qpd[r]+(QPDATA_DOUBLES-6)+lane):0.0;
            if(lane<6) ((real*)&w.qp)[QPHOT_HEAD+lane]=tailv;
            wbc_mbar_wait(barH,parH); parH^=1u;
#endif
        }
#if defined(WBC_PROFILE)
        if(PHASE!=2){ for(int k=0;k<8;k++) c.prof[k]=0; c.flops=0.0; }      // a resumed solve keeps what it had counted when it was parked
        const long long tstart=clock64();
#endif
        WBC_SYNC();
        int nWSR=nWSRmax;
        const int rcv=qp_init<true>(w,c,nWSR);
        if(PHASE==1&&rcv==QP_SUSPENDED){
#if defined(WBC_PROFILE)
            c.prof[6]=clock64()-tstart;
#endif
            park_store(w,c,park,parkCtl,parkList,r,lane);     // nothing of this robot is published yet; first[r] stays set
            while(!gang_all(1)){}
            continue;
        }
        WBC_SYNC();
        // epilogue of kernel 2 for a cold-started robot
        QPState& S=qs[r];
        const bool haveSol=(c.status==QPS_AUXILIARYQPSOLVED||c.status==QPS_HOMOTOPYQPSOLVED||c.status==QPS_SOLVED);
        LV_FOR(i,sl,42){ LV_LD(lbi,lb,i,sl); LV_LD(ubi,ub,i,sl); LV_WAIT(); LV_USE(lbi); LV_USE(ubi);
            if(LV_IN(i,42)){ S.x[i]=w.x[i]; S.lb[i]=lbi; S.ub[i]=ubi; S.bstat[i]=w.bstat[i]; if(haveSol) out.x[i*N+r]=w.x[i]; else w.x[i]=(real)out.x[i*N+r]; out.ws[i*N+r]=w.bstat[i]; } }
        LANE_FOR(i,112) S.y[i]=w.y[i];
        LV_FOR(i,sl,70){ LV_LD(al,Ax_l,i,sl); LV_LD(au,Ax_u,i,sl); LV_WAIT(); LV_USE(al); LV_USE(au);
            if(LV_IN(i,70)){ S.Ax_l[i]=al; S.Ax_u[i]=au; S.cstat[i]=w.cstat[i]; out.ws[(42+i)*N+r]=w.cstat[i]; } }
        WBC_SYNC();
        if(lane<12){ const int i=lane; const QPHotR& q=WQP;
            real s=0; for(int j=0;j<12;j++) s+=q.Mjj[12*i+j]*w.x[6+j];
            real t=0; for(int j=0;j<12;j++) t+=(q.cmask[j/3]*q.Jc[18*j+6+i])*w.x[18+j];
            out.tau[i*N+r]= (rcv==0)? (double)(s+(real)c.qg->biasj[i]-t) : 0.0;
            out.grf[i*N+r]= w.x[18+i]; }
        if(lane==0){ S.rampOffset=c.rampOffset; S.status=c.status; out.rc[r]=rcv; out.nwsr[r]=nWSR; first[r]=0;
                     if(stats){ stats[r]+=1; if(rcv!=0) stats[N+r]+=1; stats[2*N+r]+=nWSR; if(nWSR>stats[3*N+r]) stats[3*N+r]=nWSR; }
#if defined(WBC_PROFILE)
                     if(prof){ c.prof[6]=(PHASE==2?c.prof[6]:0)+(clock64()-tstart); for(int k=0;k<8;k++) prof[10*(size_t)r+k]=c.prof[k]; prof[10*(size_t)r+8]=(long long)c.flops; prof[10*(size_t)r+9]=1; }
#endif
        }
        WBC_SYNC();
        if(cohort){ while(!gang_all(1)){} }
    }
#if WBC_LV_TMEM
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if((threadIdx.x>>5)==0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(*tmBaseS),"n"(WBC_TM_ALLOC_COLS) : "memory");
#endif
}

// ---------------------------------------------------------------- kernel 3 (torque recovery, wbc.cpp:553-577) is the epilogue of kernel 2:
// x and the compact QP are still in shared memory there; a separate kernel re-read the 5.7 KB AoS QPData record per thread.
// gather of the dual solution (QProblem::getDualSolution) from the AoS solver state into the SoA output layout: inspection only
__global__ void wbc_gather_duals_kernel(int N,const QPState* qs,double* y){
    const int r=blockIdx.x*blockDim.x+threadIdx.x;
    if(r>=N) return;
    for(int k=0;k<112;k++) y[(size_t)k*N+r]=qs[r].y[k];
}

// ---------------------------------------------------------------- batched MotionPlanner (wbc_planner.cuh), one planner per thread
// mode 0: one planner period + hand-over of a published message (referenceCallback, wbc.cpp:150-185); mode 1: constructor state
struct RefOut { double *pose,*vel,*acc,*swp,*swv,*swa; unsigned char *contacts,*prevc,*switching,*published; };
__global__ void wbc_planner_kernel(int N,PlannerParams P,PlannerState* ps,const double* cmd,RefOut o,int mode,int first,int count){
    const int r=first+blockIdx.x*blockDim.x+threadIdx.x;
    if(r>=first+count||r>=N) return;
    PlannerState& s=ps[r];
    if(mode==1){ planner_init(P,s); o.published[r]=0; return; }
    for(int k=0;k<3;k++) s.cmd[k]=cmd[(size_t)k*N+r];             // input_callback (motion_planner.cpp:122-127), delivered by the spinOnce before this period
    const bool pub=planner_tick(P,s);
    o.published[r]=pub?1:0;
    if(!pub) return;
    for(int k=0;k<6;k++){ o.pose[(size_t)k*N+r]=s.pose[k]; o.vel[(size_t)k*N+r]=s.vel[k]; o.acc[(size_t)k*N+r]=s.acc[k]; }
    for(int k=0;k<12;k++){ o.swp[(size_t)k*N+r]=s.swp[k]; o.swv[(size_t)k*N+r]=s.swv[k]; o.swa[(size_t)k*N+r]=s.swa[k]; }
    bool sw=false; for(int l=0;l<4;l++){ const unsigned char c=s.contacts[l]; if(o.prevc[(size_t)l*N+r]!=c) sw=true; o.prevc[(size_t)l*N+r]=c; o.contacts[(size_t)l*N+r]=c; }
    o.switching[r]=sw?1:0;
}

// ---------------------------------------------------------------- ideal-actuator plant, one robot per thread
// The QP's equality rows are the floating-base dynamics and the stance constraints, so the generalized acceleration implied by its
// solution, nu_dot = T^-1 (a* - Tdot nu) with a* = x[0:18], is what the rigid-body model does under (tau*, f*).  T^-1 = [[X,-B],[0,1]]
// (wbc_assembly.cuh).  Semi-implicit Euler at the control rate: velocities first, then positions; the orientation is advanced by the
// world-frame rotation vector omega*dt.  A robot whose solve failed coasts (zero acceleration).
__global__ void wbc_plant_step_kernel(int N,double dt,const double* aux,const double* x,const int* rc,double* q,double* qd,double* pos,double* quat,double* twist){
    const int r=blockIdx.x*blockDim.x+threadIdx.x;
    if(r>=N) return;
    const bool ok=rc[r]==0;
    double d[3],ab[6],qdd[12],nb[6];
    for(int k=0;k<3;k++) d[k]=aux[(size_t)k*N+r];
    for(int k=0;k<6;k++) ab[k]=x[(size_t)k*N+r]-aux[(size_t)(75+k)*N+r];
    for(int k=0;k<12;k++) qdd[k]=ok?x[(size_t)(6+k)*N+r]:0.0;
    double c[3]; cross3(d,ab+3,c);
    for(int k=0;k<3;k++){ nb[k]=ab[k]+c[k]; nb[3+k]=ab[3+k]; }
    for(int i=0;i<6;i++){ double v=nb[i]; for(int j=0;j<12;j++) v-=aux[(size_t)(3+12*i+j)*N+r]*qdd[j]; nb[i]=ok?v:0.0; }
    double tw[6]; for(int k=0;k<6;k++){ tw[k]=twist[(size_t)k*N+r]+nb[k]*dt; twist[(size_t)k*N+r]=tw[k]; }
    for(int k=0;k<12;k++){ const double v=qd[(size_t)k*N+r]+qdd[k]*dt; qd[(size_t)k*N+r]=v; q[(size_t)k*N+r]+=v*dt; }
    for(int k=0;k<3;k++) pos[(size_t)k*N+r]+=tw[k]*dt;
    const double rx=tw[3]*dt, ry=tw[4]*dt, rz=tw[5]*dt, ang=sqrt(rx*rx+ry*ry+rz*rz);
    double dw=1.0,dx=0.5*rx,dy=0.5*ry,dz=0.5*rz;                      // small-angle limit of (cos(a/2), sin(a/2) r/a)
    if(ang>1e-12){ const double sh=sin(0.5*ang)/ang; dw=cos(0.5*ang); dx=sh*rx; dy=sh*ry; dz=sh*rz; }
    const double qw=quat[r],qx=quat[(size_t)N+r],qy=quat[(size_t)2*N+r],qz=quat[(size_t)3*N+r];
    double nw=dw*qw-dx*qx-dy*qy-dz*qz, nx=dw*qx+dx*qw+dy*qz-dz*qy, ny=dw*qy-dx*qz+dy*qw+dz*qx, nz=dw*qz+dx*qy-dy*qx+dz*qw;
    const double nn=sqrt(nw*nw+nx*nx+ny*ny+nz*nz); nw/=nn; nx/=nn; ny/=nn; nz/=nn;
    quat[r]=nw; quat[(size_t)N+r]=nx; quat[(size_t)2*N+r]=ny; quat[(size_t)3*N+r]=nz;
}

// wbc.cpp:176-184: isSwitchingFootState_ = (contacts != previous message's contacts); also rolls prev <- current
__global__ void wbc_derive_switching_kernel(int N,const unsigned char* contacts,unsigned char* prev,unsigned char* switching,int derive){
    const int r=blockIdx.x*blockDim.x+threadIdx.x;
    if(r>=N) return;
    bool sw=false; for(int l=0;l<4;l++){ unsigned char c=contacts[l*N+r]?1:0; if(prev[l*N+r]!=c) sw=true; prev[l*N+r]=c; }
    if(derive) switching[r]=sw?1:0;
}
// setInitialState (wbc.cpp:65-120) for robots [first, first+count)
__global__ void wbc_reset_kernel(int N,Params prm,int first,int count,double* q,double* qd,double* pos,double* quat,double* twist,double* pose,double* vel,double* acc,
                                 double* swp,double* swv,double* swa,unsigned char* contacts,unsigned char* prevc,unsigned char* switching,
                                 unsigned char* firstflag,int* rc,int* nwsr,double* tau,double* grf,double* x){
    const int r=first+blockIdx.x*blockDim.x+threadIdx.x;
    if(r>=first+count||r>=N) return;
    const double q0[12]={0.0,-0.4,0.8, 0.0,0.4,-0.8, 0.0,0.4,-0.8, 0.0,-0.4,0.8};
    for(int k=0;k<12;k++){ q[k*N+r]=q0[k]; qd[k*N+r]=0; swp[k*N+r]=0; swv[k*N+r]=0; swa[k*N+r]=0; tau[k*N+r]=0; grf[k*N+r]=0; }
    for(int k=0;k<6;k++){ twist[k*N+r]=0; pose[k*N+r]=prm.refPose[k]; vel[k*N+r]=0; acc[k*N+r]=0; }
    pos[r]=0; pos[N+r]=0; pos[2*N+r]=0.60; quat[r]=1; quat[N+r]=0; quat[2*N+r]=0; quat[3*N+r]=0;
    for(int l=0;l<4;l++){ contacts[l*N+r]=1; prevc[l*N+r]=1; }
    for(int k=0;k<42;k++) x[k*N+r]=0;                               // qpSolution_.setZero() (wbc.cpp:58)
    switching[r]=0; firstflag[r]=1; rc[r]=0; nwsr[r]=0;
}

// ================================================================= host side: C ABI (include/wbc_b200.h)
struct wbc_handle {
    int N=0, device=0; cudaStream_t stream=nullptr; bool ownStream=false;
    wbc_params params; Params devParams;
    // inputs (device, SoA)
    double *q=nullptr,*qd=nullptr,*pos=nullptr,*quat=nullptr,*twist=nullptr,*pose=nullptr,*vel=nullptr,*acc=nullptr,*swp=nullptr,*swv=nullptr,*swa=nullptr;
    unsigned char *contacts=nullptr,*prevc=nullptr,*switching=nullptr,*first=nullptr;
    // persistent per-robot state and intermediates
    LegOut* legs=nullptr; CtrlState* ctrl=nullptr; QPData* qpd=nullptr; QPState* qps=nullptr; int* counter=nullptr; long long* prof=nullptr;
    // outputs
    double *tau=nullptr,*grf=nullptr,*x=nullptr,*com=nullptr; int *rc=nullptr,*nwsr=nullptr; signed char* ws=nullptr;
    cudaEvent_t ev[4]={nullptr,nullptr,nullptr,nullptr};
    // planner / plant / rollout
    PlannerState* planner=nullptr; PlannerParams plannerParams; double* plannerCmd=nullptr; unsigned char* published=nullptr;
    double* plantAux=nullptr; int* stats=nullptr; long long tick=0;
    int smCount=0, qpBlocksPerSM=1, cohortMode=-1, updBlocksPerSM=1; size_t qpSmem=0, updSmem=0;
    double* duals=nullptr;
    QPPre* pre=nullptr; bool split=true, maybeWarm=false, setupCanary=false; size_t setupPad=0;            // two-kernel warm start (kernel 2a); maybeWarm: some robot may hold a previous solution
    // closed-loop rollout as a CUDA graph: one planner period (planner tick + graphPeriod control ticks incl. the plant) captured once, replayed per period
    cudaGraph_t graph=nullptr; cudaGraphExec_t graphExec=nullptr; int graphPeriod=0; long long graphLaunches=0; bool capturing=false, useGraph=true;
    // parking of cold solves (wbc_solve_cold_kernel): 40 KB per robot of the chunk in flight, allocated at the first all-cold solve
    QPPark* park=nullptr; int* parkList=nullptr; int* parkCtl=nullptr; int parkCap=0, parkChunk=WBC_PARK_CHUNK; bool parkMode=true;
    void* staticScratch=nullptr;                    // device address of g_wbc_static_scratch on this handle's device (wbc_solve_qp_static)
    long long launches=0;
    std::string err;
};

#define CK(call) do{ cudaError_t e_=(call); if(e_!=cudaSuccess){ h->err=std::string(#call)+": "+cudaGetErrorString(e_); return WBC_ERR_CUDA; } }while(0)

static bool is_device_ptr(const void* p){
    cudaPointerAttributes a; if(cudaPointerGetAttributes(&a,p)!=cudaSuccess){ cudaGetLastError(); return false; }
    return a.type==cudaMemoryTypeDevice||a.type==cudaMemoryTypeManaged;
}
static int copy_in(wbc_handle* h,void* dst,const void* src,size_t bytes){
    if(!src) return WBC_OK;
    CK(cudaMemcpyAsync(dst,src,bytes,is_device_ptr(src)?cudaMemcpyDeviceToDevice:cudaMemcpyHostToDevice,h->stream));
    return WBC_OK;
}
static int copy_out(wbc_handle* h,void* dst,const void* src,size_t bytes,bool& hostSeen){
    if(!dst) return WBC_OK;
    const bool dev=is_device_ptr(dst); if(!dev) hostSeen=true;
    CK(cudaMemcpyAsync(dst,src,bytes,dev?cudaMemcpyDeviceToDevice:cudaMemcpyDeviceToHost,h->stream));
    return WBC_OK;
}

extern "C" void wbc_default_params(wbc_params* p){
    if(!p) return;
    p->friction=1.0; p->loopRate=400.0; p->maxTorque=80.0; p->kpValue=6000.0; p->kpValueZ=10000.0; p->kdValue=1800.0; p->kiValue=0.0;
    p->kpSwingValue=250.0; p->kdSwingValue=20.0; p->slackWeight=1000.0;
    const double r[6]={0.0,0.0,0.5,0.0,0.0,0.0}; for(int i=0;i<6;i++) p->initialReferencePose[i]=r[i];
}

template<class T> static cudaError_t dalloc(T*& p,size_t n){ return cudaMalloc((void**)&p,n*sizeof(T)); }

static cudaError_t wbc_static_scratch_address(int device,void** out);      // device address of the static scratch of wbc_solve_qp_static (defined with it, below)
extern "C" int wbc_create(int n,int device,const wbc_params* params,void* stream,wbc_handle** out){
    if(!out||n<=0) return WBC_ERR_ARG;
    *out=nullptr;
    wbc_handle* h=new wbc_handle; h->N=n; h->device=device;
    if(params) h->params=*params; else wbc_default_params(&h->params);
    // The solver omits qpOASES' flipping-bound and zero-curvature branches; they cannot fire while H = blkdiag(I, I, I + Jst Jst', slackWeight I) >= I
    // (wbc_qp.cuh header).  slackWeight < 1 breaks that premise (H's smallest eigenvalue becomes slackWeight), so it is refused here, loudly.
    if(!(h->params.slackWeight>=1.0)||!(h->params.loopRate>0.0)){
        fprintf(stderr,"wbc_create failed: slackWeight must be >= 1 (the solver's omitted qpOASES branches rely on H >= I) and loopRate > 0\n"); delete h; return WBC_ERR_ARG; }
    auto fail=[&](int code){ fprintf(stderr,"wbc_create failed: %s\n",h->err.c_str()); wbc_destroy(h); return code; };
    cudaError_t e=cudaSetDevice(device);
    if(e!=cudaSuccess){ h->err=std::string("cudaSetDevice: ")+cudaGetErrorString(e)+" (no CPU fallback exists)"; int c=WBC_ERR_CUDA; fprintf(stderr,"wbc_create failed: %s\n",h->err.c_str()); delete h; return c; }
    cudaDeviceProp prop; if(cudaGetDeviceProperties(&prop,device)!=cudaSuccess){ h->err="cudaGetDeviceProperties failed"; return fail(WBC_ERR_CUDA); }
    h->smCount=prop.multiProcessorCount;
    if(stream){ h->stream=(cudaStream_t)stream; h->ownStream=false; }
    else { if(cudaStreamCreateWithFlags(&h->stream,cudaStreamNonBlocking)!=cudaSuccess){ h->err="cudaStreamCreate failed"; return fail(WBC_ERR_CUDA); } h->ownStream=true; }
    const size_t N=(size_t)n; bool ok=true;
    ok=ok&&dalloc(h->q,12*N)==cudaSuccess&&dalloc(h->qd,12*N)==cudaSuccess&&dalloc(h->pos,3*N)==cudaSuccess&&dalloc(h->quat,4*N)==cudaSuccess&&dalloc(h->twist,6*N)==cudaSuccess;
    ok=ok&&dalloc(h->pose,6*N)==cudaSuccess&&dalloc(h->vel,6*N)==cudaSuccess&&dalloc(h->acc,6*N)==cudaSuccess&&dalloc(h->swp,12*N)==cudaSuccess&&dalloc(h->swv,12*N)==cudaSuccess&&dalloc(h->swa,12*N)==cudaSuccess;
    ok=ok&&dalloc(h->contacts,4*N)==cudaSuccess&&dalloc(h->prevc,4*N)==cudaSuccess&&dalloc(h->switching,N)==cudaSuccess&&dalloc(h->first,N)==cudaSuccess;
    ok=ok&&dalloc(h->legs,4*N)==cudaSuccess&&dalloc(h->ctrl,N)==cudaSuccess&&dalloc(h->qpd,N)==cudaSuccess&&dalloc(h->qps,N)==cudaSuccess&&dalloc(h->counter,1)==cudaSuccess;
    ok=ok&&dalloc(h->tau,12*N)==cudaSuccess&&dalloc(h->grf,12*N)==cudaSuccess&&dalloc(h->x,42*N)==cudaSuccess&&dalloc(h->com,3*N)==cudaSuccess;
    ok=ok&&dalloc(h->rc,N)==cudaSuccess&&dalloc(h->nwsr,N)==cudaSuccess&&dalloc(h->ws,112*N)==cudaSuccess;
    if(!ok){ h->err=std::string("cudaMalloc: ")+cudaGetErrorString(cudaGetLastError()); return fail(WBC_ERR_NOMEM); }
    if(cudaMemsetAsync(h->ctrl,0,N*sizeof(CtrlState),h->stream)!=cudaSuccess||cudaMemsetAsync(h->qpd,0,N*sizeof(QPData),h->stream)!=cudaSuccess
       ||cudaMemsetAsync(h->qps,0,N*sizeof(QPState),h->stream)!=cudaSuccess||cudaMemsetAsync(h->ws,0,112*N,h->stream)!=cudaSuccess
       ||cudaMemsetAsync(h->com,0,3*N*sizeof(double),h->stream)!=cudaSuccess){ h->err=std::string("cudaMemsetAsync: ")+cudaGetErrorString(cudaGetLastError()); return fail(WBC_ERR_CUDA); }
    // constants: 13-body model table and gains
    ModelConst mc; memcpy(mc.restR,ANYMAL_restR,sizeof(mc.restR)); memcpy(mc.restP,ANYMAL_restP,sizeof(mc.restP)); memcpy(mc.axis,ANYMAL_axis,sizeof(mc.axis));
    memcpy(mc.mass,ANYMAL_mass,sizeof(mc.mass)); memcpy(mc.mcom,ANYMAL_mcom,sizeof(mc.mcom)); memcpy(mc.inertia,ANYMAL_inertia,sizeof(mc.inertia));
    memcpy(mc.footR,ANYMAL_footR,sizeof(mc.footR)); memcpy(mc.footP,ANYMAL_footP,sizeof(mc.footP));
    mc.total_mass=0; for(int i=0;i<13;i++) mc.total_mass+=mc.mass[i];                 // Model::getTotalMass (wbc.cpp:72)
    Params P; P.friction=h->params.friction; P.loopRate=h->params.loopRate; P.maxTorque=h->params.maxTorque; P.kpValue=h->params.kpValue; P.kpValueZ=h->params.kpValueZ;
    P.kdValue=h->params.kdValue; P.kiValue=h->params.kiValue; P.kpSwingValue=h->params.kpSwingValue; P.kdSwingValue=h->params.kdSwingValue; P.slackWeight=h->params.slackWeight;
    for(int i=0;i<6;i++) P.refPose[i]=h->params.initialReferencePose[i];
    h->devParams=P;
    if(cudaMemcpyToSymbol(c_model,&mc,sizeof(mc))!=cudaSuccess){
        h->err=std::string("cudaMemcpyToSymbol: ")+cudaGetErrorString(cudaGetLastError())+" (library built for sm_100a only)"; return fail(WBC_ERR_CUDA); }
    wbc_static_scratch_address(device,&h->staticScratch);                           // set-up of wbc_solve_qp_static (a failure is reported by that call, not here)
    h->updSmem=sizeof(ModelConst)+UPD_GROUPS*sizeof(UpdSmem);
    if(cudaFuncSetAttribute(wbc_update_state_kernel,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)h->updSmem)!=cudaSuccess
       ||cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->updBlocksPerSM,wbc_update_state_kernel,32,h->updSmem)!=cudaSuccess||h->updBlocksPerSM<1){
        h->err=std::string("update-state kernel geometry: ")+cudaGetErrorString(cudaGetLastError()); return fail(WBC_ERR_CUDA); }
    // solver launch geometry: one warp per CTA, as many CTAs per SM as shared memory allows
    h->qpSmem=sizeof(QPWork)*WBC_GANG+16*WBC_GANG+(WBC_LV_TMEM?16:0);                                  // + two mbarriers per warp
    if(cudaFuncSetAttribute(wbc_solve_qp_kernel,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)h->qpSmem)!=cudaSuccess
       ||cudaFuncSetAttribute(wbc_solve_cold_kernel<1>,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)h->qpSmem)!=cudaSuccess
       ||cudaFuncSetAttribute(wbc_solve_cold_kernel<2>,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)h->qpSmem)!=cudaSuccess){
        h->err=std::string("cudaFuncSetAttribute: ")+cudaGetErrorString(cudaGetLastError()); return fail(WBC_ERR_CUDA); }
    int nb=0; if(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb,wbc_solve_qp_kernel,WBC_QP_THREADS,h->qpSmem)!=cudaSuccess||nb<1){
        h->err=std::string("occupancy query: ")+cudaGetErrorString(cudaGetLastError()); return fail(WBC_ERR_CUDA); }
    h->qpBlocksPerSM=nb;
#if defined(WBC_DEV_KNOBS)                                                                                        // measurement builds only (tools/): never in the shipped library
    if(const char* e=getenv("WBC_QP_SETUP_PAD")){ int v=atoi(e); if(v>0&&v<100000) h->setupPad=(size_t)v; }          // extra shared memory per CTA of kernel 2a = fewer robots per SM (occupancy scan)
#endif
    if(cudaFuncSetAttribute(wbc_warm_setup_kernel,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)((h->setupCanary?sizeof(QPWork):QPWORK_TQ_BYTES)+h->setupPad))!=cudaSuccess){
        h->err=std::string("cudaFuncSetAttribute: ")+cudaGetErrorString(cudaGetLastError()); return fail(WBC_ERR_CUDA); }
    for(int i=0;i<4;i++) if(cudaEventCreate(&h->ev[i])!=cudaSuccess){ h->err="cudaEventCreate failed"; return fail(WBC_ERR_CUDA); }
    *out=h;
    int rcode=wbc_reset(h,0,n);
    if(rcode!=WBC_OK){ *out=nullptr; return fail(rcode); }
    return WBC_OK;
}

static void drop_graph(wbc_handle* h);
extern "C" int wbc_destroy(wbc_handle* h){
    if(!h) return WBC_ERR_ARG;
    cudaSetDevice(h->device);
    if(h->stream) cudaStreamSynchronize(h->stream);
    drop_graph(h);
    void* ptrs[]={h->q,h->qd,h->pos,h->quat,h->twist,h->pose,h->vel,h->acc,h->swp,h->swv,h->swa,h->contacts,h->prevc,h->switching,h->first,h->ctrl,h->qpd,h->qps,
                  h->counter,h->tau,h->grf,h->x,h->com,h->rc,h->nwsr,h->ws,h->prof,h->planner,h->plannerCmd,h->published,h->plantAux,h->stats,h->pre,h->duals,h->legs,h->park,h->parkList,h->parkCtl};
    for(void* p:ptrs) if(p) cudaFree(p);
    for(int i=0;i<4;i++) if(h->ev[i]) cudaEventDestroy(h->ev[i]);
    if(h->ownStream&&h->stream) cudaStreamDestroy(h->stream);
    delete h; return WBC_OK;
}

static BatchIn batch_in(const wbc_handle* h){ BatchIn b; b.q=h->q; b.qd=h->qd; b.pos=h->pos; b.quat=h->quat; b.twist=h->twist; b.pose=h->pose; b.vel=h->vel; b.acc=h->acc;
    b.swp=h->swp; b.swv=h->swv; b.swa=h->swa; b.contacts=h->contacts; b.switching=h->switching; return b; }
static BatchOut batch_out(const wbc_handle* h){ BatchOut o; o.tau=h->tau; o.grf=h->grf; o.x=h->x; o.com=h->com; o.rc=h->rc; o.nwsr=h->nwsr; o.ws=h->ws; return o; }

extern "C" int wbc_reset(wbc_handle* h,int first,int count){
    if(!h) return WBC_ERR_ARG;
    if(first<0||count<0||first+count>h->N){ h->err="wbc_reset: robot range outside [0, n_robots)"; return WBC_ERR_ARG; }
    if(count==0) return WBC_OK;
    CK(cudaSetDevice(h->device));
    wbc_reset_kernel<<<(count+127)/128,128,0,h->stream>>>(h->N,h->devParams,first,count,h->q,h->qd,h->pos,h->quat,h->twist,h->pose,h->vel,h->acc,h->swp,h->swv,h->swa,
                                                         h->contacts,h->prevc,h->switching,h->first,h->rc,h->nwsr,h->tau,h->grf,h->x);
    h->launches++; CK(cudaGetLastError());
    if(first==0&&count==h->N){ h->tick=0; h->maybeWarm=false; }
    return WBC_OK;
}
extern "C" int wbc_set_state(wbc_handle* h,const double* q,const double* qd,const double* pos,const double* quat,const double* twist){
    if(!h) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    const size_t N=(size_t)h->N, D=sizeof(double); int r;
    if((r=copy_in(h,h->q,q,12*N*D))||(r=copy_in(h,h->qd,qd,12*N*D))||(r=copy_in(h,h->pos,pos,3*N*D))||(r=copy_in(h,h->quat,quat,4*N*D))||(r=copy_in(h,h->twist,twist,6*N*D))) return r;
    return WBC_OK;
}
extern "C" int wbc_set_reference(wbc_handle* h,const double* pose,const double* vel,const double* acc,const double* swp,const double* swv,const double* swa,
                                 const unsigned char* contacts,const unsigned char* switching){
    if(!h) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    const size_t N=(size_t)h->N, D=sizeof(double); int r;
    if((r=copy_in(h,h->pose,pose,6*N*D))||(r=copy_in(h,h->vel,vel,6*N*D))||(r=copy_in(h,h->acc,acc,6*N*D))||(r=copy_in(h,h->swp,swp,12*N*D))||(r=copy_in(h,h->swv,swv,12*N*D))
       ||(r=copy_in(h,h->swa,swa,12*N*D))||(r=copy_in(h,h->contacts,contacts,4*N))||(r=copy_in(h,h->switching,switching,N))) return r;
    if(contacts){
        wbc_derive_switching_kernel<<<(h->N+127)/128,128,0,h->stream>>>(h->N,h->contacts,h->prevc,h->switching,switching?0:1);
        h->launches++; CK(cudaGetLastError());
    }
    return WBC_OK;
}
extern "C" int wbc_update_state(wbc_handle* h){
    if(!h) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    int grid=h->smCount*h->updBlocksPerSM; if(grid>(h->N+UPD_GROUPS-1)/UPD_GROUPS) grid=(h->N+UPD_GROUPS-1)/UPD_GROUPS;
    if(!h->capturing) CK(cudaEventRecord(h->ev[0],h->stream));
    wbc_rbd_legs_kernel<<<(4*h->N+127)/128,128,0,h->stream>>>(h->N,batch_in(h),h->legs);
    h->launches++; CK(cudaGetLastError());
    wbc_update_state_kernel<<<grid,32,h->updSmem,h->stream>>>(h->N,h->devParams,batch_in(h),h->legs,h->ctrl,h->qpd,h->first,h->com,h->plantAux);
    h->launches++; CK(cudaGetLastError());
    if(!h->capturing) CK(cudaEventRecord(h->ev[1],h->stream));
    return WBC_OK;
}
// ---------------------------------------------------------------- no-allocation variant of solveQP (no counterpart in the reference)
// wbc_solve_qp allocates the hand-over buffer between kernels 2a and 2 (QPPre, 22 KB per robot) with cudaMalloc at the first hot-started tick.  Callers that
// cannot allocate (CUDA graph capture, tight serving loops) use wbc_solve_qp_static: the same two kernels with that scratch -- and the work counter -- in this
// file-scope array, for batches of at most WBC_STATIC_MAX_ROBOTS (include/wbc_b200.h: 1024) robots.  Every GPU that loads the library holds the array (22.5 MB) until exit; every call on a
// device shares it, so calls that overlap on the GPU race.  Nothing in it has to start at zero except the work counter, which each call clears: kernel 2a
// rewrites all of a robot's record that kernel 2 goes on to use (the bulk copy also moves stale T entries beyond tri(nAC), which are never read -- exactly
// as with the uninitialised cudaMalloc buffer of wbc_solve_qp).
constexpr size_t WBC_STATIC_PRE_BYTES=sizeof(QPPre)*(size_t)WBC_STATIC_MAX_ROBOTS;
static __device__ __align__(256) unsigned char g_wbc_static_scratch[WBC_STATIC_PRE_BYTES+256];      // [QPPre x 1024 | work counter]
// In host code the array's name is only a stand-in: the device address is asked for once per device, with that device current, and cached here.
constexpr int WBC_STATIC_MAX_DEVICES=64;
static void* g_wbc_static_addr[WBC_STATIC_MAX_DEVICES]={};
static std::mutex g_wbc_static_mutex;
static cudaError_t wbc_static_scratch_address(int device,void** out){            // the caller has made `device` current
    *out=nullptr;
    if(device<0||device>=WBC_STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_wbc_static_mutex);
    if(!g_wbc_static_addr[device]){
        void* p=nullptr; const cudaError_t e=cudaGetSymbolAddress(&p,g_wbc_static_scratch);
        if(e!=cudaSuccess){ cudaGetLastError(); return e; }
        g_wbc_static_addr[device]=p;
    }
    *out=g_wbc_static_addr[device];
    return cudaSuccess;
}
extern "C" int wbc_solve_qp_static(wbc_handle* h){
    if(!h) return WBC_ERR_ARG;
    if(h->N>WBC_STATIC_MAX_ROBOTS){
        h->err="wbc_solve_qp_static: "+std::to_string(h->N)+" robots exceed the static scratch of "+std::to_string(WBC_STATIC_MAX_ROBOTS)+" robots (WBC_STATIC_MAX_ROBOTS); use wbc_solve_qp";
        return WBC_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    void* base=h->staticScratch;                                                  // resolved at set-up (wbc_create)
    if(!base){ const cudaError_t e=wbc_static_scratch_address(h->device,&base);
               if(e!=cudaSuccess||!base){ h->err=std::string("wbc_solve_qp_static: cudaGetSymbolAddress: ")+cudaGetErrorString(e); return WBC_ERR_CUDA; }
               h->staticScratch=base; }
    QPPre* pre=reinterpret_cast<QPPre*>(base);
    int* counter=reinterpret_cast<int*>(reinterpret_cast<unsigned char*>(base)+WBC_STATIC_PRE_BYTES);
    int grid=h->smCount*h->qpBlocksPerSM;
    if(grid>(h->N+WBC_GANG-1)/WBC_GANG) grid=(h->N+WBC_GANG-1)/WBC_GANG;
    const int cohort= h->cohortMode<0 ? (h->maybeWarm?0:1) : h->cohortMode;
    CK(cudaMemsetAsync(counter,0,sizeof(int),h->stream));                         // the one thing that must start at zero
    if(!h->capturing) CK(cudaEventRecord(h->ev[1],h->stream));
    const bool two=h->maybeWarm;                                                  // hot-started robots: kernel 2a rebuilds their factorisation into the static scratch
    if(two){
        wbc_warm_setup_kernel<<<h->N,32,(h->setupCanary?sizeof(QPWork):QPWORK_TQ_BYTES)+h->setupPad,h->stream>>>(h->N,h->qpd,h->qps,h->first,pre,h->setupCanary?1:0);
        h->launches++; CK(cudaGetLastError());
    }
    wbc_solve_qp_kernel<<<grid,WBC_QP_THREADS,h->qpSmem,h->stream>>>(h->N,h->qpd,h->qps,h->first,batch_out(h),counter,100 /* nWSR, wbc.cpp:517 */,h->prof,cohort,two?pre:nullptr,h->stats);
    h->launches++; CK(cudaGetLastError());
    h->maybeWarm=true;
    if(!h->capturing) CK(cudaEventRecord(h->ev[2],h->stream));
    return WBC_OK;
}
extern "C" int wbc_solve_qp(wbc_handle* h){
    if(!h) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    int grid=h->smCount*h->qpBlocksPerSM;
    if(grid>(h->N+WBC_GANG-1)/WBC_GANG) grid=(h->N+WBC_GANG-1)/WBC_GANG;
    // cohort scheduling only when every robot of the batch cold-starts (nothing solved since the last full reset)
    const int cohort= h->cohortMode<0 ? (h->maybeWarm?0:1) : h->cohortMode;
    CK(cudaMemsetAsync(h->counter,0,sizeof(int),h->stream));
    if(!h->capturing) CK(cudaEventRecord(h->ev[1],h->stream));
    bool two=h->split&&h->maybeWarm;
    if(two&&!h->pre&&cudaMalloc(&h->pre,sizeof(QPPre)*(size_t)h->N)!=cudaSuccess){      // 22 KB per robot; without it the solve kernel rebuilds the factorisation itself
        cudaGetLastError(); h->pre=nullptr; h->split=false; two=false;
        h->err="warm-start workspace (22 KB per robot) could not be allocated: the TQ rebuild runs inside the solve kernel (same results, slower); see wbc_warm_start_split"; }
    if(two){
        wbc_warm_setup_kernel<<<h->N,32,(h->setupCanary?sizeof(QPWork):QPWORK_TQ_BYTES)+h->setupPad,h->stream>>>(h->N,h->qpd,h->qps,h->first,h->pre,h->setupCanary?1:0);
        h->launches++; CK(cudaGetLastError());
    }
    // an all-cold batch in gang mode: two launches of wbc_solve_cold_kernel per chunk of robots (gang fetch with parking, then per-warp fetch over the parked solves);
    // bit-identical to the single launch of kernel 2, without the idle warps at the end of every gang's round
    bool parking=cohort==1&&!h->maybeWarm&&h->parkMode;
    if(parking&&!h->park){
        const int cap= h->N<h->parkChunk ? h->N : h->parkChunk;
        if(cudaMalloc(&h->park,sizeof(QPPark)*(size_t)cap)!=cudaSuccess||cudaMalloc(&h->parkList,sizeof(int)*(size_t)cap)!=cudaSuccess||cudaMalloc(&h->parkCtl,4*sizeof(int))!=cudaSuccess){
            cudaGetLastError(); if(h->park) cudaFree(h->park); if(h->parkList) cudaFree(h->parkList); if(h->parkCtl) cudaFree(h->parkCtl);
            h->park=nullptr; h->parkList=nullptr; h->parkCtl=nullptr; h->parkMode=false; parking=false;
            h->err="parking buffer (40 KB per robot of a chunk) could not be allocated: cold batches run as one launch of the general kernel (same results, slower)"; }
        else h->parkCap=cap;
    }
    if(parking){
        for(int base=0;base<h->N;base+=h->parkCap){
            const int end= base+h->parkCap<h->N ? base+h->parkCap : h->N;
            int g1=h->smCount*h->qpBlocksPerSM; if(g1>(end-base+WBC_GANG-1)/WBC_GANG) g1=(end-base+WBC_GANG-1)/WBC_GANG;
            if(base>0) CK(cudaMemsetAsync(h->counter,0,sizeof(int),h->stream));
            CK(cudaMemsetAsync(h->parkCtl,0,4*sizeof(int),h->stream));
            wbc_solve_cold_kernel<1><<<g1,WBC_QP_THREADS,h->qpSmem,h->stream>>>(h->N,h->qpd,h->qps,h->first,batch_out(h),h->counter,100 /* nWSR, wbc.cpp:517 */,h->prof,h->stats,base,end,h->park,h->parkCtl,h->parkList);
            h->launches++; CK(cudaGetLastError());
            wbc_solve_cold_kernel<2><<<g1,WBC_QP_THREADS,h->qpSmem,h->stream>>>(h->N,h->qpd,h->qps,h->first,batch_out(h),h->parkCtl+1,100,h->prof,h->stats,base,end,h->park,h->parkCtl,h->parkList);
            h->launches++; CK(cudaGetLastError());
        }
    } else {
    wbc_solve_qp_kernel<<<grid,WBC_QP_THREADS,h->qpSmem,h->stream>>>(h->N,h->qpd,h->qps,h->first,batch_out(h),h->counter,100 /* nWSR, wbc.cpp:517 */,h->prof,cohort,two?h->pre:nullptr,h->stats);
    h->launches++; CK(cudaGetLastError());
    }
    h->maybeWarm=true;
    if(!h->capturing) CK(cudaEventRecord(h->ev[2],h->stream));
    return WBC_OK;
}
extern "C" int wbc_compute_joint_torques(wbc_handle* h){
    if(!h) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    // tau and f* were written by the epilogue of wbc_solve_qp_kernel from the solution in shared memory; nothing is left to launch
    if(!h->capturing) CK(cudaEventRecord(h->ev[3],h->stream));
    return WBC_OK;
}
extern "C" int wbc_step(wbc_handle* h){ int r; if((r=wbc_update_state(h))||(r=wbc_solve_qp(h))||(r=wbc_compute_joint_torques(h))) return r; return WBC_OK; }
extern "C" int wbc_synchronize(wbc_handle* h){ if(!h) return WBC_ERR_ARG; CK(cudaSetDevice(h->device)); CK(cudaStreamSynchronize(h->stream)); return WBC_OK; }

extern "C" int wbc_get_outputs(wbc_handle* h,double* tau,double* grf,double* x,int* rc,int* nwsr,signed char* ws,double* com){
    if(!h) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    const size_t N=(size_t)h->N, D=sizeof(double); bool host=false; int r;
    if((r=copy_out(h,tau,h->tau,12*N*D,host))||(r=copy_out(h,grf,h->grf,12*N*D,host))||(r=copy_out(h,x,h->x,42*N*D,host))||(r=copy_out(h,rc,h->rc,N*sizeof(int),host))
       ||(r=copy_out(h,nwsr,h->nwsr,N*sizeof(int),host))||(r=copy_out(h,ws,h->ws,112*N,host))||(r=copy_out(h,com,h->com,3*N*D,host))) return r;
    if(host) CK(cudaStreamSynchronize(h->stream));
    return WBC_OK;
}
extern "C" int wbc_get_duals(wbc_handle* h,double* y){
    if(!h||!y){ if(h) h->err="wbc_get_duals: null argument"; return WBC_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    const size_t N=(size_t)h->N;
    if(!h->duals){ if(dalloc(h->duals,112*N)!=cudaSuccess){ h->err="cudaMalloc (duals)"; cudaGetLastError(); return WBC_ERR_NOMEM; } }
    wbc_gather_duals_kernel<<<(h->N+127)/128,128,0,h->stream>>>(h->N,h->qps,h->duals); h->launches++; CK(cudaGetLastError());
    bool host=false; int r; if((r=copy_out(h,y,h->duals,112*N*sizeof(double),host))) return r;
    if(host) CK(cudaStreamSynchronize(h->stream));
    return WBC_OK;
}
extern "C" int wbc_get_dense_qp(wbc_handle* h,int robot,double* H,double* g,double* A,double* lbA,double* ubA){
    if(!h) return WBC_ERR_ARG;
    if(robot<0||robot>=h->N){ h->err="wbc_get_dense_qp: robot index outside [0, n_robots)"; return WBC_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    QPData q; CK(cudaMemcpyAsync(&q,h->qpd+robot,sizeof(QPData),cudaMemcpyDeviceToHost,h->stream)); CK(cudaStreamSynchronize(h->stream));
    if(H) for(int i=0;i<42;i++) for(int j=0;j<42;j++) H[i*42+j]=Hel(q,i,j);
    if(g) for(int i=0;i<42;i++) g[i]=(i>=18&&i<30)?q.gf[i-18]:0.0;
    if(A) for(int i=0;i<70;i++) for(int j=0;j<42;j++) A[i*42+j]=Ael(q,i,j);
    if(lbA) memcpy(lbA,q.lbA,sizeof(q.lbA)); if(ubA) memcpy(ubA,q.ubA,sizeof(q.ubA));
    return WBC_OK;
}

// ================================================================= planner, plant, rollout (include/wbc_b200.h, section "next rows")
static void fill_quintic(double* a,double T,double vi,double vf){       // generateQuinticPolynomial (motion_planner.cpp:77-97), host doubles
    const double T2=T*T, T3=T2*T, T4=T3*T, T5=T4*T;
    a[0]=0.0; a[1]=vi; a[2]=0.0; a[3]=(10.0 - 4.0 * vf * T - 6.0 * vi * T) / T3; a[4]=(-15.0 + 7.0 * vf * T + 8.0 * vi * T) / T4; a[5]=(6.0 - 3.0 * vf * T - 3.0 * vi * T) / T5;
}
extern "C" void wbc_planner_default_params(wbc_planner_params* p){
    if(!p) return;
    p->step_length=0.1; p->height_control_point=0.1; p->x_offset=0.50; p->y_offset=0.33; p->step_duration=0.2; p->body_height=0.50; p->body_final_velocity=0.40; p->dt=0.01;
}
static RefOut ref_out(const wbc_handle* h){ RefOut o; o.pose=h->pose; o.vel=h->vel; o.acc=h->acc; o.swp=h->swp; o.swv=h->swv; o.swa=h->swa; o.contacts=h->contacts; o.prevc=h->prevc;
    o.switching=h->switching; o.published=h->published; return o; }
extern "C" int wbc_planner_reset(wbc_handle* h,int first,int count){
    if(!h) return WBC_ERR_ARG;
    if(!h->planner||first<0||count<0||first+count>h->N){ h->err="wbc_planner_reset: needs wbc_planner_create and a robot range inside [0, n_robots)"; return WBC_ERR_ARG; }
    if(count==0) return WBC_OK;
    CK(cudaSetDevice(h->device));
    wbc_planner_kernel<<<(count+127)/128,128,0,h->stream>>>(h->N,h->plannerParams,h->planner,h->plannerCmd,ref_out(h),1,first,count);
    h->launches++; CK(cudaGetLastError());
    return WBC_OK;
}
extern "C" int wbc_planner_create(wbc_handle* h,const wbc_planner_params* p){
    if(h) drop_graph(h);
    if(!h) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    wbc_planner_params d; if(p) d=*p; else wbc_planner_default_params(&d);
    PlannerParams& P=h->plannerParams;
    P.step_length=d.step_length; P.height_control_point=d.height_control_point; P.x_offset=d.x_offset; P.y_offset=d.y_offset; P.step_duration=d.step_duration;
    P.cycle_duration=4*d.step_duration; P.body_height=d.body_height; P.body_final_velocity=d.body_final_velocity; P.dt=d.dt;     // load_parameters (motion_planner.cpp:99-119)
    fill_quintic(P.poly_foot,P.step_duration,0.0,0.0); fill_quintic(P.poly_body_start,P.cycle_duration,0.0,P.body_final_velocity);
    fill_quintic(P.poly_body_continue,P.cycle_duration,P.body_final_velocity,P.body_final_velocity);                                  // motion_planner.cpp:190-195
    const size_t N=(size_t)h->N;
    if(!h->planner){
        if(dalloc(h->planner,N)!=cudaSuccess||dalloc(h->plannerCmd,3*N)!=cudaSuccess||dalloc(h->published,N)!=cudaSuccess){ h->err="cudaMalloc (planner)"; return WBC_ERR_NOMEM; }
        CK(cudaMemsetAsync(h->plannerCmd,0,3*N*sizeof(double),h->stream));
    }
    return wbc_planner_reset(h,0,h->N);
}
extern "C" int wbc_planner_set_cmd_vel(wbc_handle* h,const double* vx,const double* vy,const double* yaw_rate){
    if(!h) return WBC_ERR_ARG;
    if(!h->planner){ h->err="planner call before wbc_planner_create"; return WBC_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    const size_t N=(size_t)h->N, D=sizeof(double); int r;
    if((r=copy_in(h,h->plannerCmd,vx,N*D))||(r=copy_in(h,h->plannerCmd+N,vy,N*D))||(r=copy_in(h,h->plannerCmd+2*N,yaw_rate,N*D))) return r;
    return WBC_OK;
}
extern "C" int wbc_planner_tick(wbc_handle* h){
    if(!h) return WBC_ERR_ARG;
    if(!h->planner){ h->err="planner call before wbc_planner_create"; return WBC_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    wbc_planner_kernel<<<(h->N+127)/128,128,0,h->stream>>>(h->N,h->plannerParams,h->planner,h->plannerCmd,ref_out(h),0,0,h->N);
    h->launches++; CK(cudaGetLastError());
    return WBC_OK;
}
extern "C" int wbc_get_reference(wbc_handle* h,double* pose,double* vel,double* acc,double* swp,double* swv,double* swa,unsigned char* contacts,unsigned char* switching,
                                 unsigned char* published){
    if(!h) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    const size_t N=(size_t)h->N, D=sizeof(double); bool host=false; int r;
    if((r=copy_out(h,pose,h->pose,6*N*D,host))||(r=copy_out(h,vel,h->vel,6*N*D,host))||(r=copy_out(h,acc,h->acc,6*N*D,host))||(r=copy_out(h,swp,h->swp,12*N*D,host))
       ||(r=copy_out(h,swv,h->swv,12*N*D,host))||(r=copy_out(h,swa,h->swa,12*N*D,host))||(r=copy_out(h,contacts,h->contacts,4*N,host))||(r=copy_out(h,switching,h->switching,N,host))) return r;
    if(published){ if(!h->published) return WBC_ERR_ARG; if((r=copy_out(h,published,h->published,N,host))) return r; }
    if(host) CK(cudaStreamSynchronize(h->stream));
    return WBC_OK;
}
extern "C" int wbc_get_state(wbc_handle* h,double* q,double* qd,double* pos,double* quat,double* twist){
    if(!h) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    const size_t N=(size_t)h->N, D=sizeof(double); bool host=false; int r;
    if((r=copy_out(h,q,h->q,12*N*D,host))||(r=copy_out(h,qd,h->qd,12*N*D,host))||(r=copy_out(h,pos,h->pos,3*N*D,host))||(r=copy_out(h,quat,h->quat,4*N*D,host))
       ||(r=copy_out(h,twist,h->twist,6*N*D,host))) return r;
    if(host) CK(cudaStreamSynchronize(h->stream));
    return WBC_OK;
}
extern "C" int wbc_plant_enable(wbc_handle* h){
    if(!h) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    const size_t N=(size_t)h->N;
    if(!h->plantAux){ if(dalloc(h->plantAux,81*N)!=cudaSuccess){ h->err="cudaMalloc (plant)"; return WBC_ERR_NOMEM; } CK(cudaMemsetAsync(h->plantAux,0,81*N*sizeof(double),h->stream)); }
    if(!h->stats){ if(dalloc(h->stats,4*N)!=cudaSuccess){ h->err="cudaMalloc (stats)"; return WBC_ERR_NOMEM; } CK(cudaMemsetAsync(h->stats,0,4*N*sizeof(int),h->stream)); }
    return WBC_OK;
}
extern "C" int wbc_plant_step(wbc_handle* h){
    if(!h) return WBC_ERR_ARG;
    if(!h->plantAux){ h->err="wbc_plant_step before wbc_plant_enable"; return WBC_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    wbc_plant_step_kernel<<<(h->N+127)/128,128,0,h->stream>>>(h->N,1.0/h->devParams.loopRate,h->plantAux,h->x,h->rc,h->q,h->qd,h->pos,h->quat,h->twist);
    h->launches++; CK(cudaGetLastError());
    return WBC_OK;
}
static void drop_graph(wbc_handle* h){
    if(h->graphExec){ cudaGraphExecDestroy(h->graphExec); h->graphExec=nullptr; }
    if(h->graph){ cudaGraphDestroy(h->graph); h->graph=nullptr; }
    h->graphPeriod=0;
}
static int rollout_tick(wbc_handle* h,int planner_every){
    int r;
    if(h->tick%planner_every==0) if((r=wbc_planner_tick(h))) return r;
    if((r=wbc_step(h))||(r=wbc_plant_step(h))) return r;
    h->tick++;
    return WBC_OK;
}
// One planner period of the steady-state closed loop (every robot hot-started, kernel 2a's workspace allocated) is the same 1 + 6 x planner_every
// stream operations every time: it is captured into a CUDA graph once and replayed, so a small batch (<= 888 robots: 0.2 ms ticks) pays one graph
// launch per period instead of 25 kernel launches.  Results are those of the plain loop (same kernels, same order); the per-kernel event timers
// (wbc_last_kernel_ms) are not updated by replayed periods.
static int build_rollout_graph(wbc_handle* h,int planner_every){
    drop_graph(h);
    const long long tick0=h->tick, l0=h->launches;
    if(cudaStreamBeginCapture(h->stream,cudaStreamCaptureModeThreadLocal)!=cudaSuccess){ cudaGetLastError(); return WBC_ERR_CUDA; }
    h->capturing=true; int r=WBC_OK;
    for(int t=0;t<planner_every&&r==WBC_OK;t++) r=rollout_tick(h,planner_every);
    h->capturing=false;
    cudaGraph_t g=nullptr; const cudaError_t e=cudaStreamEndCapture(h->stream,&g);
    h->tick=tick0; h->graphLaunches=h->launches-l0; h->launches=l0;                  // nothing ran yet
    if(r!=WBC_OK||e!=cudaSuccess||!g){ if(g) cudaGraphDestroy(g); cudaGetLastError(); return r!=WBC_OK?r:WBC_ERR_CUDA; }
    h->graph=g;
    if(cudaGraphInstantiate(&h->graphExec,h->graph,0)!=cudaSuccess){ cudaGetLastError(); drop_graph(h); return WBC_ERR_CUDA; }
    h->graphPeriod=planner_every;
    return WBC_OK;
}
extern "C" int wbc_rollout(wbc_handle* h,int ticks,int planner_every){
    if(!h) return WBC_ERR_ARG;
    if(ticks<0||planner_every<1||!h->plantAux||!h->planner){ h->err="wbc_rollout: needs wbc_planner_create, wbc_plant_enable, ticks >= 0 and planner_every >= 1"; return WBC_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    int r, t=0;
    while(t<ticks){
        const bool steady=h->useGraph&&h->maybeWarm&&h->split&&h->pre&&h->cohortMode<=0&&!h->prof&&h->tick%planner_every==0&&ticks-t>=planner_every;
        if(steady&&(h->graphExec==nullptr||h->graphPeriod!=planner_every)){
            if(build_rollout_graph(h,planner_every)!=WBC_OK){ h->useGraph=false; h->err="wbc_rollout: CUDA graph capture failed, continuing with plain launches"; continue; } }
        if(steady){
            CK(cudaGraphLaunch(h->graphExec,h->stream));
            h->tick+=planner_every; t+=planner_every; h->launches+=h->graphLaunches;
        } else { if((r=rollout_tick(h,planner_every))) return r; t++; }
    }
    return WBC_OK;
}
extern "C" int wbc_rollout_stats(wbc_handle* h,int* ticks,int* failed_ticks,int* nwsr_sum,int* nwsr_max,int reset){
    if(!h) return WBC_ERR_ARG;
    if(!h->stats){ h->err="wbc_rollout_stats before wbc_plant_enable"; return WBC_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    const size_t N=(size_t)h->N; bool host=false; int r;
    if((r=copy_out(h,ticks,h->stats,N*sizeof(int),host))||(r=copy_out(h,failed_ticks,h->stats+N,N*sizeof(int),host))||(r=copy_out(h,nwsr_sum,h->stats+2*N,N*sizeof(int),host))
       ||(r=copy_out(h,nwsr_max,h->stats+3*N,N*sizeof(int),host))) return r;
    if(host) CK(cudaStreamSynchronize(h->stream));
    if(reset) CK(cudaMemsetAsync(h->stats,0,4*N*sizeof(int),h->stream));
    return WBC_OK;
}
// exact build (-DWBC_EXACT): [0] full linear-independence tests run, [1] of them inside the undecided band of the product build's |Z'a| shortcut,
// [2] tests whose verdict the shortcut would have got wrong (must stay 0).  Other builds: WBC_ERR_ARG.
extern "C" int wbc_debug_counters(wbc_handle* h,unsigned long long* out3,int reset){
    if(!h||!out3) return WBC_ERR_ARG;
#if defined(WBC_EXACT)
    CK(cudaSetDevice(h->device)); CK(cudaStreamSynchronize(h->stream));
    unsigned long long v[4]={0,0,0,0}; CK(cudaMemcpyFromSymbol(v,g_wbc_li,sizeof(v)));
    out3[0]=v[0]; out3[1]=v[1]; out3[2]=v[2];
    if(reset){ const unsigned long long z[4]={0,0,0,0}; CK(cudaMemcpyToSymbol(g_wbc_li,z,sizeof(z))); }
    return WBC_OK;
#else
    (void)reset; h->err="wbc_debug_counters: only in the exact build (libwbc_b200_exact.so, -DWBC_EXACT)"; return WBC_ERR_ARG;
#endif
}
extern "C" int wbc_num_robots(const wbc_handle* h){ return h?h->N:0; }
extern "C" int wbc_warm_start_split(const wbc_handle* h){ return (h&&h->split)?1:0; }
// explicit switches (tests, measurements); every setting gives bit-identical results
extern "C" int wbc_set_option(wbc_handle* h,const char* key,int value){
    if(!h||!key) return WBC_ERR_ARG;
    const std::string k(key);
    drop_graph(h);                                                            // every switch below changes what a captured period would launch
    if(k=="rollout_graph"){ h->useGraph=value!=0; return WBC_OK; }            // 0: wbc_rollout issues plain launches every tick
    if(k=="warm_start_split"){ h->split=value!=0; return WBC_OK; }            // 0: rebuild the TQ factorisation inside the solve kernel
    if(k=="park_chunk"){                                                      // robots per pair of launches of a parked cold batch (tests; before the first cold solve)
        if(value<1||h->park){ h->err="wbc_set_option: park_chunk needs a positive value and must come before the first cold solve"; return WBC_ERR_ARG; }
        h->parkChunk=value; return WBC_OK; }
    if(k=="park"){ h->parkMode=value!=0; return WBC_OK; }                       // 0: all-cold batches run as one launch of the general kernel (no parking of solves)
    if(k=="cohort"){ h->cohortMode=value; return WBC_OK; }                    // -1 auto, 0 never, 1 always (gang scheduling of kernel 2)
    if(k=="setup_canary"){                                                    // guard the prefix allocation of kernel 2a (tests)
        h->setupCanary=value!=0;
        CK(cudaSetDevice(h->device));
        CK(cudaFuncSetAttribute(wbc_warm_setup_kernel,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)((h->setupCanary?sizeof(QPWork):QPWORK_TQ_BYTES)+h->setupPad)));
        return WBC_OK; }
    h->err="wbc_set_option: unknown key "+k; return WBC_ERR_ARG;
}
extern "C" long long wbc_launch_count(const wbc_handle* h){ return h?h->launches:0; }
extern "C" int wbc_last_kernel_ms(wbc_handle* h,float* rbd_ms,float* qp_ms,float* torque_ms){
    if(!h) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device)); CK(cudaStreamSynchronize(h->stream));
    float a=0,b=0,c=0;
    if(cudaEventElapsedTime(&a,h->ev[0],h->ev[1])!=cudaSuccess) a=-1; if(cudaEventElapsedTime(&b,h->ev[1],h->ev[2])!=cudaSuccess) b=-1;
    if(cudaEventElapsedTime(&c,h->ev[2],h->ev[3])!=cudaSuccess) c=-1; cudaGetLastError();
    if(rbd_ms) *rbd_ms=a; if(qp_ms) *qp_ms=b; if(torque_ms) *torque_ms=c;
    return WBC_OK;
}
// per-robot breakdown of the last solve (10 x int64 per robot: QPCtx::prof[0..7], algorithmic flops, cold flag); enable!=0 allocates the buffer
extern "C" int wbc_profile(wbc_handle* h,int enable,long long* host_out){
    if(!h) return WBC_ERR_ARG;
#if !defined(WBC_PROFILE)
    (void)enable; (void)host_out; h->err="wbc_profile: this is the product build; the phase timers and the flop counter exist only in libwbc_b200_prof.so (-DWBC_PROFILE)"; return WBC_ERR_ARG;
#else
    CK(cudaSetDevice(h->device));
    if(enable&&!h->prof){ CK(cudaMalloc((void**)&h->prof,10*(size_t)h->N*sizeof(long long))); CK(cudaMemsetAsync(h->prof,0,10*(size_t)h->N*sizeof(long long),h->stream)); }
    if(host_out&&h->prof){ CK(cudaMemcpyAsync(host_out,h->prof,10*(size_t)h->N*sizeof(long long),cudaMemcpyDeviceToHost,h->stream)); CK(cudaStreamSynchronize(h->stream)); }
    return WBC_OK;
#endif
}
// FP64 FMA peak of the device, measured: 8 independent DFMA chains per thread, 1024 threads x 2 CTAs per SM
__global__ void wbc_fp64_peak_kernel(double* out,int iters){
    double a0=1.0+threadIdx.x*1e-9,a1=a0+1,a2=a0+2,a3=a0+3,a4=a0+4,a5=a0+5,a6=a0+6,a7=a0+7; const double m=1.0000001,b=1e-9;
    for(int i=0;i<iters;i++){ a0=fma(a0,m,b);a1=fma(a1,m,b);a2=fma(a2,m,b);a3=fma(a3,m,b);a4=fma(a4,m,b);a5=fma(a5,m,b);a6=fma(a6,m,b);a7=fma(a7,m,b); }
    if(a0+a1+a2+a3+a4+a5+a6+a7==12345.678) out[0]=a0;
}
extern "C" int wbc_measure_fp64_peak(wbc_handle* h,double* gflops){
    if(!h||!gflops) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    const int iters=1<<16, grid=h->smCount*2, block=1024; cudaEvent_t e0,e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    wbc_fp64_peak_kernel<<<grid,block,0,h->stream>>>(h->com,256); h->launches++;
    double best=0;
    for(int rep=0;rep<3;rep++){
        CK(cudaEventRecord(e0,h->stream)); wbc_fp64_peak_kernel<<<grid,block,0,h->stream>>>(h->com,iters); h->launches++; CK(cudaEventRecord(e1,h->stream));
        CK(cudaEventSynchronize(e1)); float ms=0; CK(cudaEventElapsedTime(&ms,e0,e1));
        const double gf=2.0*8.0*iters*(double)grid*block/(ms*1e-3)/1e9; if(gf>best) best=gf; }
    cudaEventDestroy(e0); cudaEventDestroy(e1); *gflops=best; return WBC_OK;
}
// shared-memory bandwidth of the device, measured: conflict-free 64-bit loads (what the solver issues), 1024 threads x 2 CTAs per SM
__global__ void wbc_smem_peak_kernel(unsigned long long* out,int iters){
    __shared__ unsigned long long s[4096];
    for(int i=threadIdx.x;i<4096;i+=blockDim.x) s[i]=(unsigned long long)i*0x9E3779B97F4A7C15ull;
    __syncthreads();
    unsigned long long a0=0,a1=0,a2=0,a3=0; int idx=threadIdx.x;
    #pragma unroll 4
    for(int it=0;it<iters;it++){ a0^=s[idx]; a1^=s[(idx+1024)&4095]; a2^=s[(idx+2048)&4095]; a3^=s[(idx+3072)&4095]; idx=(idx+32)&4095; }
    if((a0^a1^a2^a3)==0x1234567ull) out[0]=a0;
}
extern "C" int wbc_measure_smem_peak(wbc_handle* h,double* gbytes_per_s){
    if(!h||!gbytes_per_s) return WBC_ERR_ARG;
    CK(cudaSetDevice(h->device));
    const int iters=1<<14, grid=h->smCount*2, block=1024; cudaEvent_t e0,e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    wbc_smem_peak_kernel<<<grid,block,0,h->stream>>>((unsigned long long*)h->com,64); h->launches++;
    double best=0;
    for(int rep=0;rep<3;rep++){
        CK(cudaEventRecord(e0,h->stream)); wbc_smem_peak_kernel<<<grid,block,0,h->stream>>>((unsigned long long*)h->com,iters); h->launches++; CK(cudaEventRecord(e1,h->stream));
        CK(cudaEventSynchronize(e1)); float ms=0; CK(cudaEventElapsedTime(&ms,e0,e1));
        const double gb=8.0*4.0*iters*(double)grid*block/(ms*1e-3)/1e9; if(gb>best) best=gb; }
    cudaEventDestroy(e0); cudaEventDestroy(e1); *gbytes_per_s=best; return WBC_OK;
}
extern "C" const char* wbc_last_error(const wbc_handle* h){ return h?h->err.c_str():"null handle"; }

// ================================================================= one batch over several GPUs of the box (include/wbc_b200.h, "multi-GPU")
// The path shards by robot index with no exchange: device k owns the contiguous range [first_k, first_k + count_k) for the lifetime
// of the object -- one wbc_handle, one stream per device, all issued from the calling host thread.  Global arrays are structure-of-arrays over the
// WHOLE batch, a[j*N + i]; a shard's slice of one array is a strided block, moved with ONE cudaMemcpy2DAsync per array and device.
struct wbc_multi {
    int N=0, nDev=0;
    std::vector<wbc_handle*> h; std::vector<int> first, count, device;
    std::vector<cudaEvent_t> g0, g1;            // brackets of the last output copy per device (gather timing)
    std::string err;
};
#define MCK(call) do{ cudaError_t e_=(call); if(e_!=cudaSuccess){ m->err=std::string(#call)+": "+cudaGetErrorString(e_); return WBC_ERR_CUDA; } }while(0)

extern "C" int wbc_multi_destroy(wbc_multi* m){
    if(!m) return WBC_ERR_ARG;
    for(size_t k=0;k<m->h.size();k++){ if(m->h[k]){ cudaSetDevice(m->device[k]); if(k<m->g0.size()&&m->g0[k]) cudaEventDestroy(m->g0[k]); if(k<m->g1.size()&&m->g1[k]) cudaEventDestroy(m->g1[k]); wbc_destroy(m->h[k]); } }
    delete m; return WBC_OK;
}
extern "C" int wbc_multi_create(int n_robots,const int* device_ids,int n_devices,const wbc_params* params,wbc_multi** out){
    if(!out||n_robots<=0||n_devices<=0||!device_ids||n_devices>n_robots) return WBC_ERR_ARG;
    *out=nullptr;
    wbc_multi* m=new wbc_multi; m->N=n_robots; m->nDev=n_devices;
    const int base=n_robots/n_devices, rem=n_robots%n_devices;            // same split as workloads.shard_range: the remainder goes to the lowest shards
    for(int k=0;k<n_devices;k++){
        const int cnt=base+(k<rem?1:0), fst=k*base+(k<rem?k:rem);
        wbc_handle* h=nullptr; const int rc=wbc_create(cnt,device_ids[k],params,nullptr,&h);
        if(rc!=WBC_OK){ fprintf(stderr,"wbc_multi_create: shard %d on device %d failed with %d\n",k,device_ids[k],rc); wbc_multi_destroy(m); return rc; }
        m->h.push_back(h); m->first.push_back(fst); m->count.push_back(cnt); m->device.push_back(device_ids[k]);
        cudaEvent_t a=nullptr,b=nullptr; cudaSetDevice(device_ids[k]); cudaEventCreate(&a); cudaEventCreate(&b); m->g0.push_back(a); m->g1.push_back(b);
    }
    // direct peer copies for the gather where the topology allows them (NVLink / NVSwitch); without peer access the runtime stages through the host
    for(int a=0;a<n_devices;a++) for(int b=0;b<n_devices;b++) if(a!=b){ int can=0; cudaDeviceCanAccessPeer(&can,device_ids[a],device_ids[b]);
        if(can){ cudaSetDevice(device_ids[a]); if(cudaDeviceEnablePeerAccess(device_ids[b],0)!=cudaSuccess) cudaGetLastError(); } }
    *out=m; return WBC_OK;
}
extern "C" int wbc_multi_num_devices(const wbc_multi* m){ return m?m->nDev:0; }
extern "C" int wbc_multi_num_robots(const wbc_multi* m){ return m?m->N:0; }
extern "C" int wbc_multi_shard(const wbc_multi* m,int k,int* device,int* first,int* count){
    if(!m||k<0||k>=m->nDev) return WBC_ERR_ARG;
    if(device) *device=m->device[k]; if(first) *first=m->first[k]; if(count) *count=m->count[k]; return WBC_OK;
}
extern "C" wbc_handle* wbc_multi_handle(wbc_multi* m,int k){ return (m&&k>=0&&k<m->nDev)?m->h[k]:nullptr; }
extern "C" const char* wbc_multi_last_error(const wbc_multi* m){
    if(!m) return "null handle";
    if(!m->err.empty()) return m->err.c_str();
    for(auto* h:m->h) if(h&&!h->err.empty()) return h->err.c_str();
    return "";
}
// slice [first, first+count) of a global (rows x N) array <-> the shard's own (rows x count) array: one strided copy
static int multi_in(wbc_multi* m,int k,void* dst,const void* src,int rows,size_t elem){
    if(!src) return WBC_OK;
    const wbc_handle* h=m->h[k];
    MCK(cudaMemcpy2DAsync(dst,(size_t)m->count[k]*elem,(const char*)src+(size_t)m->first[k]*elem,(size_t)m->N*elem,(size_t)m->count[k]*elem,rows,cudaMemcpyDefault,h->stream));
    return WBC_OK;
}
static int multi_out(wbc_multi* m,int k,void* dst,const void* src,int rows,size_t elem){
    if(!dst) return WBC_OK;
    const wbc_handle* h=m->h[k];
    MCK(cudaMemcpy2DAsync((char*)dst+(size_t)m->first[k]*elem,(size_t)m->N*elem,src,(size_t)m->count[k]*elem,(size_t)m->count[k]*elem,rows,cudaMemcpyDefault,h->stream));
    return WBC_OK;
}
extern "C" int wbc_multi_reset(wbc_multi* m){ if(!m) return WBC_ERR_ARG; for(int k=0;k<m->nDev;k++){ int r=wbc_reset(m->h[k],0,m->count[k]); if(r) return r; } return WBC_OK; }
extern "C" int wbc_multi_set_state(wbc_multi* m,const double* q,const double* qd,const double* pos,const double* quat,const double* twist){
    if(!m) return WBC_ERR_ARG;
    for(int k=0;k<m->nDev;k++){ wbc_handle* h=m->h[k]; MCK(cudaSetDevice(h->device)); int r; const size_t D=sizeof(double);
        if((r=multi_in(m,k,h->q,q,12,D))||(r=multi_in(m,k,h->qd,qd,12,D))||(r=multi_in(m,k,h->pos,pos,3,D))||(r=multi_in(m,k,h->quat,quat,4,D))||(r=multi_in(m,k,h->twist,twist,6,D))) return r; }
    return WBC_OK;
}
extern "C" int wbc_multi_set_reference(wbc_multi* m,const double* pose,const double* vel,const double* acc,const double* swp,const double* swv,const double* swa,
                                       const unsigned char* contacts,const unsigned char* switching){
    if(!m) return WBC_ERR_ARG;
    for(int k=0;k<m->nDev;k++){ wbc_handle* h=m->h[k]; MCK(cudaSetDevice(h->device)); int r; const size_t D=sizeof(double);
        if((r=multi_in(m,k,h->pose,pose,6,D))||(r=multi_in(m,k,h->vel,vel,6,D))||(r=multi_in(m,k,h->acc,acc,6,D))||(r=multi_in(m,k,h->swp,swp,12,D))||(r=multi_in(m,k,h->swv,swv,12,D))
           ||(r=multi_in(m,k,h->swa,swa,12,D))||(r=multi_in(m,k,h->contacts,contacts,4,1))||(r=multi_in(m,k,h->switching,switching,1,1))) return r;
        if(contacts&&(r=wbc_set_reference(h,nullptr,nullptr,nullptr,nullptr,nullptr,nullptr,h->contacts,switching?h->switching:nullptr))) return r; }   // derives / rolls the switching flag like the single-device call
    return WBC_OK;
}
extern "C" int wbc_multi_step(wbc_multi* m){ if(!m) return WBC_ERR_ARG; for(int k=0;k<m->nDev;k++){ int r=wbc_step(m->h[k]); if(r) return r; } return WBC_OK; }
extern "C" int wbc_multi_synchronize(wbc_multi* m){ if(!m) return WBC_ERR_ARG; for(int k=0;k<m->nDev;k++){ int r=wbc_synchronize(m->h[k]); if(r) return r; } return WBC_OK; }
// results of every shard into global (rows x N) arrays: host memory, or device memory of ANY device of the box -- the latter is the optional
// single gather of SURVEY 8e (about 200 B per robot over NVLink; peer copies where enabled).  Returns after all copies have completed.
extern "C" int wbc_multi_get_outputs(wbc_multi* m,double* tau,double* grf,double* x,int* rc,int* nwsr,signed char* ws,double* com){
    if(!m) return WBC_ERR_ARG;
    for(int k=0;k<m->nDev;k++){ wbc_handle* h=m->h[k]; MCK(cudaSetDevice(h->device)); int r; const size_t D=sizeof(double);
        MCK(cudaEventRecord(m->g0[k],h->stream));
        if((r=multi_out(m,k,tau,h->tau,12,D))||(r=multi_out(m,k,grf,h->grf,12,D))||(r=multi_out(m,k,x,h->x,42,D))||(r=multi_out(m,k,rc,h->rc,1,sizeof(int)))
           ||(r=multi_out(m,k,nwsr,h->nwsr,1,sizeof(int)))||(r=multi_out(m,k,ws,h->ws,112,1))||(r=multi_out(m,k,com,h->com,3,D))) return r;
        MCK(cudaEventRecord(m->g1[k],h->stream)); }
    return wbc_multi_synchronize(m);
}
// duration [ms] of the output copies of the last wbc_multi_get_outputs, max over the devices (each measured with CUDA events on its own stream)
extern "C" int wbc_multi_last_gather_ms(wbc_multi* m,float* ms){
    if(!m||!ms) return WBC_ERR_ARG;
    float worst=0; for(int k=0;k<m->nDev;k++){ MCK(cudaSetDevice(m->device[k])); float t=0; if(cudaEventElapsedTime(&t,m->g0[k],m->g1[k])!=cudaSuccess){ cudaGetLastError(); t=-1; } if(t>worst) worst=t; }
    *ms=worst; return WBC_OK;
}
