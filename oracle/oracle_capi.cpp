// ORACLE -- TEST INFRASTRUCTURE ONLY. C entry points (ctypes) over the restated controller + real qpOASES.
// Array layout mirrors the product C-ABI (include/wbc_b200.h): structure-of-arrays, instance index fastest,
// element k of instance i lives at a[k*N + i].
#include "wbc_oracle.hpp"
#include "planner_oracle.hpp"
#include <qpOASES.hpp>
#include <vector>
#include <thread>
#include <atomic>
#include <chrono>
#include <cstdlib>

using namespace oracle;

struct RobotState { double q[12],qd[12],pos[3],quat[4],twist[6]; };
struct OracleBatch {
    int N; std::vector<std::unique_ptr<WbcOracle>> inst;
    std::vector<RobotState> state;            // what the state callbacks last delivered (kept for the closed-loop rollout)
};

static Params params_from(const double* p){
    Params P;
    if(p){ P.friction=p[0]; P.loopRate=p[1]; P.maxTorque=p[2]; P.kpValue=p[3]; P.kpValueZ=p[4]; P.kdValue=p[5]; P.kiValue=p[6];
           P.kpSwingValue=p[7]; P.kdSwingValue=p[8]; P.slackWeight=p[9]; for(int i=0;i<6;i++) P.refData[i]=p[10+i]; }
    return P;
}

extern "C" {

void* orc_create(int N,int lumped,const double* params16){
    OracleBatch* b=new OracleBatch; b->N=N; b->inst.resize(N); b->state.resize(N);
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
        RobotState& S=b->state[i]; for(int k=0;k<12;k++){ S.q[k]=qq[k]; S.qd[k]=dd[k]; } for(int k=0;k<3;k++) S.pos[k]=p[k]; for(int k=0;k<4;k++) S.quat[k]=qu[k]; for(int k=0;k<6;k++) S.twist[k]=tw[k];
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

// same with an iteration limit PER QP (debugging aid: the iterate after k working-set changes of a hot start is read back after RET_MAX_NWSR_REACHED)
void orc_qp_sequence_nwsr(int T,const double* H,const double* g,const double* A,const double* lbA,const double* ubA,const int* nWSRmax,
                          double* x,double* y,signed char* ws,int* rc,int* nwsr){
    qpOASES::SQProblem qp(42,70);
    for(int t=0;t<T;t++){
        qpOASES::int_t nWSR=nWSRmax[t]; qpOASES::Options o; o.setToReliable(); o.printLevel=qpOASES::PL_LOW; qp.setOptions(o);
        int r = (t==0) ? qp.init(H+t*1764,g+t*42,A+t*2940,nullptr,nullptr,lbA+t*70,ubA+t*70,nWSR,0)
                       : qp.hotstart(H+t*1764,g+t*42,A+t*2940,nullptr,nullptr,lbA+t*70,ubA+t*70,nWSR,0);
        rc[t]=r; nwsr[t]=(int)nWSR;
        double xs[42]={0},ys[112]={0},w[112]={0};
        qp.getPrimalSolution(xs); qp.getDualSolution(ys); qp.getWorkingSet(w);
        for(int k=0;k<42;k++) x[t*42+k]=xs[k];
        for(int k=0;k<112;k++){ y[t*112+k]=ys[k]; ws[t*112+k]=(signed char)w[k]; }
    }
}

// ---- restated MotionPlanner (planner_oracle.cpp): K planner periods of ONE planner; cmd[K][3] = (linear.x, linear.y, angular.z) per period.
// out: published[K] (bytes), msg[K][54] = pose6 vel6 acc6 swPos12 swVel12 swAcc12 (the latched ref_msg_ at the end of the period), contacts[K][4]
void orc_planner_run(const double* params8,int K,const double* cmd,unsigned char* published,double* msg,unsigned char* contacts){
    PlannerParams P;
    if(params8){ P.step_length=params8[0]; P.height_control_point=params8[1]; P.x_offset=params8[2]; P.y_offset=params8[3]; P.step_duration=params8[4];
                 P.body_height=params8[5]; P.body_final_velocity=params8[6]; P.dt=params8[7]; }
    std::vector<std::array<double,3>> c(K); for(int k=0;k<K;k++) c[k]={cmd[3*k],cmd[3*k+1],cmd[3*k+2]};
    std::vector<PlannerRecord> r=planner_run(P,c);
    for(int k=0;k<K;k++){ const ReferenceMsg& m=r[k].msg; published[k]=r[k].published?1:0; double* o=msg+54*(size_t)k;
        for(int i=0;i<6;i++){ o[i]=m.desiredComPose[i]; o[6+i]=m.desiredComVelocity[i]; o[12+i]=m.desiredComAcceleration[i]; }
        for(int i=0;i<12;i++){ o[18+i]=m.desiredSwingLegsPosition[i]; o[30+i]=m.desiredSwingLegsVelocity[i]; o[42+i]=m.desiredSwingLegsAcceleration[i]; }
        for(int i=0;i<4;i++) contacts[4*(size_t)k+i]=m.footContacts[i]; }
}

// ---- closed-loop rollout: restated planner -> restated controller + real qpOASES -> plant ----
// The PLANT has no counterpart in the reference (it runs against Gazebo); it is DEFINED here, densely: nu_dot = T^-1 (a - Tdot nu) with
// the controller's own 18x18 T and Tdot of this tick and a = x[0:18], zero if the solve failed; semi-implicit Euler at loopRate;
// orientation advanced by the world-frame rotation vector omega*dt and re-normalised.  The product's closed-form version must agree.
static void plant_step(WbcOracle& w,RobotState& S){
    const double dt=1.0/w.params.loopRate;
    Mat<18,1> nu,a; for(int k=0;k<6;k++) nu(k,0)=S.twist[k]; for(int k=0;k<12;k++) nu(6+k,0)=S.qd[k];
    for(int k=0;k<18;k++) a(k,0)=w.qpSolution_[k];
    Mat<18,1> nud=inverse(w.transformationMatrix_)*(a-w.transformationMatrixDot_*nu);
    if(w.qpReturnValue_!=0) nud=Mat<18,1>();
    for(int k=0;k<6;k++) S.twist[k]+=nud(k,0)*dt;
    for(int k=0;k<12;k++){ S.qd[k]+=nud(6+k,0)*dt; S.q[k]+=S.qd[k]*dt; }
    for(int k=0;k<3;k++) S.pos[k]+=S.twist[k]*dt;
    const double rx=S.twist[3]*dt, ry=S.twist[4]*dt, rz=S.twist[5]*dt, ang=std::sqrt(rx*rx+ry*ry+rz*rz);
    double dw=1.0,dx=0.5*rx,dy=0.5*ry,dz=0.5*rz;
    if(ang>1e-12){ const double sh=std::sin(0.5*ang)/ang; dw=std::cos(0.5*ang); dx=sh*rx; dy=sh*ry; dz=sh*rz; }
    const double qw=S.quat[0],qx=S.quat[1],qy=S.quat[2],qz=S.quat[3];
    double nw=dw*qw-dx*qx-dy*qy-dz*qz, nx=dw*qx+dx*qw+dy*qz-dz*qy, ny=dw*qy-dx*qz+dy*qw+dz*qx, nz=dw*qz+dx*qy-dy*qx+dz*qw;
    const double nn=std::sqrt(nw*nw+nx*nx+ny*ny+nz*nz);
    S.quat[0]=nw/nn; S.quat[1]=nx/nn; S.quat[2]=ny/nn; S.quat[3]=nz/nn;
}
// K control ticks from the current state; tick t (counted from tick0) runs one planner period first when t % planner_every == 0.
// cmd[3*N] SoA = constant /cmd_vel per robot.  Per-tick records (any may be NULL): tau[K][12][N], rc[K][N], nwsr[K][N], q[K][12][N], pos[K][3][N].
// Call once per batch (the planner is re-created at period 0 on every call).
double orc_rollout(void* h,int K,int planner_every,const double* plannerParams8,const double* cmd,int nthreads,
                   double* tau,int* rc,int* nwsr,double* qrec,double* posrec){
    OracleBatch* b=(OracleBatch*)h; const int N=b->N;
    PlannerParams P;
    if(plannerParams8){ P.step_length=plannerParams8[0]; P.height_control_point=plannerParams8[1]; P.x_offset=plannerParams8[2]; P.y_offset=plannerParams8[3];
                        P.step_duration=plannerParams8[4]; P.body_height=plannerParams8[5]; P.body_final_velocity=plannerParams8[6]; P.dt=plannerParams8[7]; }
    const int periods=(K+planner_every-1)/planner_every;
    if(nthreads<=0) nthreads=orc_max_threads();
    if(nthreads>N) nthreads=N;
    auto t0=std::chrono::steady_clock::now();
    std::atomic<int> next(0);
    auto work=[&](){ for(;;){ int i=next.fetch_add(1); if(i>=N) break;
        WbcOracle& w=*b->inst[i]; RobotState& S=b->state[i];
        std::vector<std::array<double,3>> c(periods,std::array<double,3>{cmd[i],cmd[N+i],cmd[2*N+i]});
        std::vector<PlannerRecord> plan=planner_run(P,c);
        for(int t=0;t<K;t++){
            if(t%planner_every==0){ const PlannerRecord& r=plan[t/planner_every];
                if(r.published) w.setReference(r.msg.desiredComPose,r.msg.desiredComVelocity,r.msg.desiredComAcceleration,r.msg.desiredSwingLegsPosition,
                                               r.msg.desiredSwingLegsVelocity,r.msg.desiredSwingLegsAcceleration,r.msg.footContacts,-1); }
            w.setJointState(S.q,S.qd); w.setFloatingBaseState(S.pos,S.quat,S.twist);
            w.step();
            const size_t o=(size_t)t;
            if(tau) for(int k=0;k<12;k++) tau[(o*12+k)*N+i]=(w.qpReturnValue_==0)?w.tau[k]:0.0;
            if(rc) rc[o*N+i]=w.qpReturnValue_; if(nwsr) nwsr[o*N+i]=w.nWSR_;
            plant_step(w,S);
            if(qrec) for(int k=0;k<12;k++) qrec[(o*12+k)*N+i]=S.q[k];
            if(posrec) for(int k=0;k<3;k++) posrec[(o*3+k)*N+i]=S.pos[k];
        } } };
    std::vector<std::thread> th; for(int t=1;t<nthreads;t++) th.emplace_back(work);
    work(); for(auto& t:th) t.join();
    return std::chrono::duration<double>(std::chrono::steady_clock::now()-t0).count();
}
void orc_get_state(void* h,double* q,double* qd,double* pos,double* quat,double* twist){
    OracleBatch* b=(OracleBatch*)h; const int N=b->N;
    for(int i=0;i<N;i++){ const RobotState& S=b->state[i];
        for(int k=0;k<12;k++){ q[k*N+i]=S.q[k]; qd[k*N+i]=S.qd[k]; } for(int k=0;k<3;k++) pos[k*N+i]=S.pos[k];
        for(int k=0;k<4;k++) quat[k*N+i]=S.quat[k]; for(int k=0;k<6;k++) twist[k*N+i]=S.twist[k]; }
}

}  // extern "C"
