// ORACLE -- TEST INFRASTRUCTURE ONLY. C entry points (ctypes) over the restated controller + real qpOASES.
// Array layout mirrors the product C-ABI (include/wbc_b200.h): structure-of-arrays, instance index fastest,
// element k of instance i lives at a[k*N + i].
#include "wbc_oracle.hpp"
#include <qpOASES.hpp>
#include <vector>
#include <thread>
#include <atomic>
#include <chrono>
#include <cstdlib>

using namespace oracle;

struct OracleBatch {
    int N; std::vector<std::unique_ptr<WbcOracle>> inst;
};

static Params params_from(const double* p){
    Params P;
    if(p){ P.friction=p[0]; P.loopRate=p[1]; P.maxTorque=p[2]; P.kpValue=p[3]; P.kpValueZ=p[4]; P.kdValue=p[5]; P.kiValue=p[6];
           P.kpSwingValue=p[7]; P.kdSwingValue=p[8]; P.slackWeight=p[9]; for(int i=0;i<6;i++) P.refData[i]=p[10+i]; }
    return P;
}

extern "C" {

void* orc_create(int N,int lumped,const double* params16){
    OracleBatch* b=new OracleBatch; b->N=N; b->inst.resize(N);
    Params P=params_from(params16);
    const Tree& t=lumped?anymal_lumped_tree():anymal_full_tree();
    for(int i=0;i<N;i++) b->inst[i].reset(new WbcOracle(t,P));
    return b;
}
void orc_destroy(void* h){ delete (OracleBatch*)h; }
int orc_max_threads();
int orc_max_threads(){ int n=(int)std::thread::hardware_concurrency(); return n>0?n:1; }

void orc_set_state(void* h,const double* q,const double* qd,const double* pos,const double* quat,const double* twist){
    OracleBatch* b=(OracleBatch*)h; const int N=b->N;
    for(int i=0;i<N;i++){
        double qq[12],dd[12],p[3],qu[4],tw[6];
        for(int k=0;k<12;k++){ qq[k]=q[k*N+i]; dd[k]=qd[k*N+i]; }
        for(int k=0;k<3;k++) p[k]=pos[k*N+i];
        for(int k=0;k<4;k++) qu[k]=quat[k*N+i];
        for(int k=0;k<6;k++) tw[k]=twist[k*N+i];
        b->inst[i]->setJointState(qq,dd); b->inst[i]->setFloatingBaseState(p,qu,tw);
    }
}
void orc_set_reference(void* h,const double* pose,const double* vel,const double* acc,const double* sp,const double* sv,const double* sa,
                       const unsigned char* contacts,const unsigned char* switching){
    OracleBatch* b=(OracleBatch*)h; const int N=b->N;
    for(int i=0;i<N;i++){
        double a[6],v[6],c[6],p12[12],v12[12],a12[12]; unsigned char ct[4];
        for(int k=0;k<6;k++){ a[k]=pose[k*N+i]; v[k]=vel[k*N+i]; c[k]=acc[k*N+i]; }
        for(int k=0;k<12;k++){ p12[k]=sp[k*N+i]; v12[k]=sv[k*N+i]; a12[k]=sa[k*N+i]; }
        for(int k=0;k<4;k++) ct[k]=contacts[k*N+i];
        b->inst[i]->setReference(a,v,c,p12,v12,a12,ct,switching?(int)switching[i]:-1);
    }
}
// returns wall seconds of the parallel region; nthreads<=0 -> all
double orc_step(void* h,int nthreads,int assemble_only){
    OracleBatch* b=(OracleBatch*)h; const int N=b->N;
    if(nthreads<=0) nthreads=orc_max_threads();
    if(nthreads>N) nthreads=N;
    auto t0=std::chrono::steady_clock::now();
    std::atomic<int> next(0);
    auto work=[&](){ for(;;){ int i=next.fetch_add(1); if(i>=N) break; b->inst[i]->skipSolve=(assemble_only!=0); b->inst[i]->step(); } };
    std::vector<std::thread> th; for(int t=1;t<nthreads;t++) th.emplace_back(work);
    work(); for(auto& t:th) t.join();
    return std::chrono::duration<double>(std::chrono::steady_clock::now()-t0).count();
}
void orc_get_outputs(void* h,double* tau,double* grf,double* x,int* rc,int* nwsr,signed char* ws,double* y){
    OracleBatch* b=(OracleBatch*)h; const int N=b->N;
    for(int i=0;i<N;i++){ WbcOracle& w=*b->inst[i];
        for(int k=0;k<12;k++){ if(tau) tau[k*N+i]=w.tau[k]; if(grf) grf[k*N+i]=w.grf[k]; }
        if(x) for(int k=0;k<42;k++) x[k*N+i]=w.qpSolution_[k];
        if(rc) rc[i]=w.qpReturnValue_; if(nwsr) nwsr[i]=w.nWSR_;
        if(ws) for(int k=0;k<112;k++) ws[k*N+i]=(signed char)w.workingSet_[k];
        if(y) for(int k=0;k<112;k++) y[k*N+i]=w.dual_[k];
    }
}
// dense per-instance dumps (row-major), any pointer may be NULL
void orc_get_qp(void* h,int i,double* H,double* g,double* A,double* lbA,double* ubA){
    WbcOracle& w=*((OracleBatch*)h)->inst[i];
    if(H) memcpy(H,w.H,sizeof(w.H)); if(g) memcpy(g,w.g,sizeof(w.g)); if(A) memcpy(A,w.A,sizeof(w.A));
    if(lbA) memcpy(lbA,w.lbA,sizeof(w.lbA)); if(ubA) memcpy(ubA,w.ubA,sizeof(w.ubA));
}
void orc_get_intermediates(void* h,int i,double* M,double* T,double* Mc,double* Jst,double* Jsw,double* bias_c,double* coriolis,
                           double* com,double* vcom6,double* pose6,double* footJ,double* footPos,double* footVel,double* TdotInv,double* W){
    WbcOracle& w=*((OracleBatch*)h)->inst[i];
    if(M) memcpy(M,w.massMatrix_.a.data(),324*8); if(T) memcpy(T,w.transformationMatrix_.a.data(),324*8);
    if(Mc) memcpy(Mc,w.centroidMassMatrix_.a.data(),324*8);
    if(Jst) memcpy(Jst,w.centroidStanceJacobian_.a.data(),216*8); if(Jsw) memcpy(Jsw,w.centroidSwingJacobian_.a.data(),216*8);
    if(bias_c) memcpy(bias_c,w.centroidGeneralizedBias_.a.data(),18*8); if(coriolis) memcpy(coriolis,w.coriolisBias_.a.data(),18*8);
    if(com) memcpy(com,w.centerOfMassPosition_,3*8); if(vcom6) memcpy(vcom6,w.centerOfMassVelocity_,6*8); if(pose6) memcpy(pose6,w.currentPose_,6*8);
    if(footJ) memcpy(footJ,w.footJacobianLin_.a.data(),216*8);
    if(footPos) memcpy(footPos,w.swingFootPosition_,12*8); if(footVel) memcpy(footVel,w.swingFootVelocity_,12*8);
    if(TdotInv) memcpy(TdotInv,w.transformationMatrixDotInverse_.a.data(),324*8); if(W) memcpy(W,w.desiredWrench_,6*8);
}

// ---- rigid-body-dynamics only (identity tests; full vs lumped tree) ----
// state: pos3, quat4(wxyz), twist6 (mixed), q12, dq12 ; outputs row-major
void orc_rbd_eval(int lumped,const double* pos,const double* quat,const double* twist,const double* q,const double* dq,
                  double* M,double* J4x6x18,double* h,double* g,double* com,double* vcom,double* footPos,double* footVel6,
                  const double* acc18,double* id18){
    Params P; WbcOracle tmp(lumped?anymal_lumped_tree():anymal_full_tree(),P);   // reuse the quaternion conversion
    tmp.setFloatingBaseState(pos,quat,twist);
    KinDyn kd(lumped?anymal_lumped_tree():anymal_full_tree());
    kd.setRobotState(tmp.T_world_base_,q,twist,dq,v3(0,0,gravityAcceleration));
    V3 c=kd.getCenterOfMassPosition(), vc=kd.getCenterOfMassVelocity();
    for(int k=0;k<3;k++){ com[k]=c(k); vcom[k]=vc(k); }
    kd.getFreeFloatingMassMatrix(M);
    for(int l=0;l<4;l++){ kd.getFrameFreeFloatingJacobian(l,J4x6x18+l*108);
        Transform Hf=kd.getWorldTransform(l); Sp v=kd.getFrameVel(l);
        for(int k=0;k<3;k++){ footPos[3*l+k]=Hf.p(k); footVel6[6*l+k]=v.lin(k); footVel6[6*l+3+k]=v.ang(k); } }
    kd.generalizedBiasForces(h); kd.generalizedGravityForces(g);
    if(acc18&&id18) kd.inverseDynamicsMixed(acc18,acc18+6,id18);
}

// ---- the real qpOASES driven exactly like wbc.cpp:517-541 on a sequence of dense QPs (solver parity tests) ----
// QPs are stored back to back: H[T][42*42], g[T][42], A[T][70*42], lbA[T][70], ubA[T][70]; tick 0 = init, then hotstart.
void orc_qp_sequence(int T,const double* H,const double* g,const double* A,const double* lbA,const double* ubA,int nWSRmax,
                     double* x,double* y,signed char* ws,int* rc,int* nwsr){
    qpOASES::SQProblem qp(42,70);
    for(int t=0;t<T;t++){
        qpOASES::int_t nWSR=nWSRmax; qpOASES::Options o; o.setToReliable(); o.printLevel=getenv("ORC_PRINT")?qpOASES::PL_MEDIUM:qpOASES::PL_LOW; qp.setOptions(o);
        int r = (t==0) ? qp.init(H+t*1764,g+t*42,A+t*2940,nullptr,nullptr,lbA+t*70,ubA+t*70,nWSR,0)
                       : qp.hotstart(H+t*1764,g+t*42,A+t*2940,nullptr,nullptr,lbA+t*70,ubA+t*70,nWSR,0);
        rc[t]=r; nwsr[t]=(int)nWSR;
        double xs[42]={0},ys[112]={0},w[112]={0};
        qp.getPrimalSolution(xs); qp.getDualSolution(ys); qp.getWorkingSet(w);
        for(int k=0;k<42;k++) x[t*42+k]=xs[k];
        for(int k=0;k<112;k++){ y[t*112+k]=ys[k]; ws[t*112+k]=(signed char)w[k]; }
    }
}

}  // extern "C"
