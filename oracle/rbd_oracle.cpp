// ORACLE -- TEST INFRASTRUCTURE ONLY (see rbd_oracle.hpp header). Restates iDynTree 12.3.2.
#include "rbd_oracle.hpp"
#include "anymal_tree_full.h"
#include "../quadrupedwholebodycontroller_b200/csrc/anymal_model.h"

namespace oracle {

const Tree& anymal_full_tree(){
    static Tree t;
    if(!t.n){ t.n=ANYMAL_FULL_NLINKS; t.parent=ANYMAL_FULL_parent; t.dof=ANYMAL_FULL_dof; t.restR=ANYMAL_FULL_restR; t.restP=ANYMAL_FULL_restP;
        t.axis=ANYMAL_FULL_axis; t.mass=ANYMAL_FULL_mass; t.mcom=ANYMAL_FULL_mcom; t.inertia=ANYMAL_FULL_inertia;
        t.foot_link=ANYMAL_FULL_foot_link; t.footR=ANYMAL_FULL_footR; t.footP=ANYMAL_FULL_footP; t.total_mass=0;
        // Model::getTotalMass sums link masses (lib/idyntree/src/model/src/Model.cpp getTotalMass)
        for(int i=0;i<t.n;i++) t.total_mass+=t.mass[i]; }
    return t;
}
const Tree& anymal_lumped_tree(){
    static Tree t;
    if(!t.n){ t.n=ANYMAL_NLINKS; t.parent=ANYMAL_parent; t.dof=ANYMAL_dof; t.restR=ANYMAL_restR; t.restP=ANYMAL_restP;
        t.axis=ANYMAL_axis; t.mass=ANYMAL_mass; t.mcom=ANYMAL_mcom; t.inertia=ANYMAL_inertia;
        t.foot_link=ANYMAL_foot_link; t.footR=ANYMAL_footR; t.footP=ANYMAL_footP; t.total_mass=0;
        for(int i=0;i<t.n;i++) t.total_mass+=t.mass[i]; }
    return t;
}

static inline M3 m3_from(const double* p){ M3 r; std::memcpy(r.m,p,9*sizeof(double)); return r; }
static inline V3 v3_from(const double* p){ return v3(p[0],p[1],p[2]); }
static inline Inertia link_inertia(const Tree&t,int i){ Inertia I; I.m=t.mass[i]; I.mc=v3_from(t.mcom+3*i); I.I=m3_from(t.inertia+9*i); return I; }
// RevoluteJoint.cpp:189-196 S_link2_link1: axis in the child frame, origin on the child origin => [0; axis]
static inline Sp link_S(const Tree&t,int i){ Sp s; s.lin=v3(0,0,0); s.ang=v3_from(t.axis+3*i); return s; }

KinDyn::KinDyn(const Tree& t): tree_(t), q_(12,0.0), dq_(12,0.0), linkPos_(t.n), pXc_(t.n), linkVel_(t.n), crb_(t.n) {
    world_H_base_=tf_identity(); baseVelBody_=sp_zero(); gravityBase_=v3(0,0,0); totalMomentum_=sp_zero();
    std::memset(rawM_,0,sizeof(rawM_));
}

// RevoluteJoint.cpp:128-134: link1_X_link2 = axis_wrt_link1.getRotationTransform(q) * rest
// with axis_wrt_link1 = rest * axis_child (:235)  ==  rest * Rot(axis_child, q) because the axis passes
// through the child origin. FixedJoint: the rest transform.
Transform KinDyn::jointTransform(int i) const {
    Transform rest; rest.R=m3_from(tree_.restR+9*i); rest.p=v3_from(tree_.restP+3*i);
    if(tree_.dof[i]<0) return rest;
    Transform rot; rot.R=rot_axis(v3_from(tree_.axis+3*i), q_[tree_.dof[i]]); rot.p=v3(0,0,0);
    return rest*rot;
}

// KinDynComputations.cpp:785-839 (MIXED branch :822-825; gravity :813-815)
void KinDyn::setRobotState(const Transform& wHb,const double* q,const double* baseVelMixed,const double* dq,const V3& gravity){
    fkOk_=false; massOk_=false;
    world_H_base_=wHb;
    for(int i=0;i<12;i++){ q_[i]=q[i]; dq_[i]=dq[i]; }
    M3 bRw=transpose(wHb.R);
    gravityBase_=bRw*gravity;
    baseVelBody_.lin=bRw*v3(baseVelMixed[0],baseVelMixed[1],baseVelMixed[2]);
    baseVelBody_.ang=bRw*v3(baseVelMixed[3],baseVelMixed[4],baseVelMixed[5]);
    for(int i=1;i<tree_.n;i++) pXc_[i]=jointTransform(i);
}

// ForwardKinematics.cpp:144-187 + RevoluteJoint.cpp:265-282 (computeChildVel)
void KinDyn::computeFwdKinematics(){
    if(fkOk_) return;
    for(int i=0;i<tree_.n;i++){
        int p=tree_.parent[i];
        if(p<0){ linkPos_[i]=world_H_base_; linkVel_[i]=baseVelBody_; continue; }
        linkPos_[i]=linkPos_[p]*pXc_[i];
        Sp v=xform_motion(inverse(pXc_[i]),linkVel_[p]);
        if(tree_.dof[i]>=0) v=v+dq_[tree_.dof[i]]*link_S(tree_,i);
        linkVel_[i]=v;
    }
    fkOk_=true;
}

// KinDynComputations.cpp:390-418 -> CompositeRigidBodyAlgorithm (Dynamics.cpp:146-273) and
// ComputeLinearAndAngularMomentum (Dynamics.cpp:26-42)
void KinDyn::computeRawMassMatrixAndTotalMomentum(){
    if(massOk_) return;
    const int n=tree_.n;
    std::memset(rawM_,0,sizeof(rawM_));
    for(int i=0;i<n;i++) crb_[i]=link_inertia(tree_,i);
    for(int i=n-1;i>=1;i--){
        int p=tree_.parent[i];
        crb_[p]=crb_[p]+xform_inertia(pXc_[i],crb_[i]);
        if(tree_.dof[i]<0) continue;
        Sp S=link_S(tree_,i);
        Sp F=crb_[i]*S;
        int di=tree_.dof[i];
        rawM_[(6+di)*18+6+di]=dot(S,F);
        int a=i;
        while(tree_.parent[tree_.parent[a]]>=0){            // while lambda(j) != base
            F=xform_force(pXc_[a],F);
            a=tree_.parent[a];
            if(tree_.dof[a]>=0){
                int dj=tree_.dof[a];
                double h=dot(link_S(tree_,a),F);
                rawM_[(6+di)*18+6+dj]=h; rawM_[(6+dj)*18+6+di]=h;
            }
        }
        F=xform_force(pXc_[a],F);                            // into the base frame
        double f6[6]={F.lin(0),F.lin(1),F.lin(2),F.ang(0),F.ang(1),F.ang(2)};
        for(int k=0;k<6;k++){ rawM_[k*18+6+di]=f6[k]; rawM_[(6+di)*18+k]=f6[k]; }
    }
    // locked inertia (SpatialInertia::asMatrix): [[m 1, -S(mc)],[S(mc), I]]
    { const Inertia&B=crb_[0]; M3 S=skew(B.mc);
      for(int r=0;r<3;r++)for(int c=0;c<3;c++){
        rawM_[r*18+c]=(r==c)?B.m:0.0; rawM_[r*18+3+c]=-S(r,c); rawM_[(3+r)*18+c]=S(r,c); rawM_[(3+r)*18+3+c]=B.I(r,c); } }
    computeFwdKinematics();
    totalMomentum_=sp_zero();
    for(int i=0;i<n;i++) totalMomentum_=totalMomentum_+xform_force(linkPos_[i],link_inertia(tree_,i)*linkVel_[i]);
    massOk_=true;
}

// KinDynComputations.cpp:1915-1924
V3 KinDyn::getCenterOfMassPosition(){
    computeRawMassMatrixAndTotalMomentum();
    const Inertia&B=crb_[0];
    V3 c = (std::fabs(B.m)>0) ? v3(B.mc(0)/B.m,B.mc(1)/B.m,B.mc(2)/B.m) : v3(0,0,0);
    return world_H_base_.R*c+world_H_base_.p;
}
// KinDynComputations.cpp:1941-1958
V3 KinDyn::getCenterOfMassVelocity(){
    computeRawMassMatrixAndTotalMomentum();
    V3 com=getCenterOfMassPosition();
    Transform X; X.R=m3_identity(); X.p=-com;
    Sp h=xform_force(X,totalMomentum_);
    double m=crb_[0].m;
    return v3(h.lin(0)/m,h.lin(1)/m,h.lin(2)/m);
}

// KinDynComputations.cpp:2606-2630; right side :2074-2102 (cols 0..5 times Ad(base_R_world)),
// left side :2130-2152 (rows 0..5 times Ad*(world_R_base)); rotation-only adjoints are blkdiag(R,R)
void KinDyn::getFreeFloatingMassMatrix(double* M){
    computeRawMassMatrixAndTotalMomentum();
    std::memcpy(M,rawM_,sizeof(rawM_));
    const M3&R=world_H_base_.R;
    for(int r=0;r<18;r++) for(int b=0;b<2;b++){
        double t[3]; for(int c=0;c<3;c++){ double s=0; for(int k=0;k<3;k++) s+=M[r*18+3*b+k]*R(c,k); t[c]=s; }   // row * R^T
        for(int c=0;c<3;c++) M[r*18+3*b+c]=t[c];
    }
    for(int c=0;c<18;c++) for(int b=0;b<2;b++){
        double t[3]; for(int r=0;r<3;r++){ double s=0; for(int k=0;k<3;k++) s+=R(r,k)*M[(3*b+k)*18+c]; t[r]=s; }
        for(int r=0;r<3;r++) M[(3*b+r)*18+c]=t[r];
    }
}

// KinDynComputations.cpp:1299-1340 (feet are links in the full tree, fixed frames of SHANK in the lumped one)
Transform KinDyn::getWorldTransform(int foot){
    computeFwdKinematics();
    Transform lHf; lHf.R=m3_from(tree_.footR+9*foot); lHf.p=v3_from(tree_.footP+3*foot);
    return linkPos_[tree_.foot_link[foot]]*lHf;
}

// KinDynComputations.cpp:1409-1445 MIXED: world_R_frame * (frame_X_link * v_link)
Sp KinDyn::getFrameVel(int foot){
    computeFwdKinematics();
    Transform lHf; lHf.R=m3_from(tree_.footR+9*foot); lHf.p=v3_from(tree_.footP+3*foot);
    Sp vb=xform_motion(inverse(lHf),linkVel_[tree_.foot_link[foot]]);
    Transform wHf=getWorldTransform(foot);
    Sp r; r.lin=wHf.R*vb.lin; r.ang=wHf.R*vb.ang; return r;
}

// KinDynComputations.cpp:1607-1679 (MIXED: jacobFrame_X_world=(1,-p_frame), baseFrame_X_jacobBaseFrame=(base_R_world,0))
// + FreeFloatingJacobianUsingLinkPos (Jacobians.cpp:16-52)
void KinDyn::getFrameFreeFloatingJacobian(int foot,double* J){
    computeFwdKinematics();
    std::memset(J,0,6*18*sizeof(double));
    Transform wHf=getWorldTransform(foot);
    Transform jXw; jXw.R=m3_identity(); jXw.p=-wHf.p;
    Transform bXj; bXj.R=transpose(world_H_base_.R); bXj.p=v3(0,0,0);
    Transform X=jXw*linkPos_[0]*bXj;
    M3 SpR=skew(X.p)*X.R;                                  // asAdjointTransform = [[R, S(p)R],[0,R]]
    for(int r=0;r<3;r++)for(int c=0;c<3;c++){ J[r*18+c]=X.R(r,c); J[r*18+3+c]=SpR(r,c); J[(3+r)*18+3+c]=X.R(r,c); }
    int l=tree_.foot_link[foot];
    while(l!=0){
        if(tree_.dof[l]>=0){
            Sp s=xform_motion(jXw,xform_motion(linkPos_[l],link_S(tree_,l)));
            int c=6+tree_.dof[l];
            for(int k=0;k<3;k++){ J[k*18+c]=s.lin(k); J[(3+k)*18+c]=s.ang(k); }
        }
        l=tree_.parent[l];
    }
}

// ForwardAccKinematics (ForwardKinematics.cpp:189-226, RevoluteJoint.cpp:315-326) + RNEADynamicPhase
// (Dynamics.cpp:65-144, torque RevoluteJoint.cpp:342-352); base wrench rotated to MIXED (KDC.cpp:2653-2668)
void KinDyn::rnea(const Sp& baseProperAcc,const double* ddq,bool zeroVel,double* out){
    computeFwdKinematics();
    const int n=tree_.n;
    std::vector<Sp> acc(n), f(n);
    for(int i=0;i<n;i++){
        int p=tree_.parent[i];
        if(p<0){ acc[i]=baseProperAcc; continue; }
        Sp a=xform_motion(inverse(pXc_[i]),acc[p]);
        if(tree_.dof[i]>=0){
            Sp S=link_S(tree_,i);
            a=a+(ddq?ddq[tree_.dof[i]]:0.0)*S;
            if(!zeroVel) a=a+cross_motion(linkVel_[i],dq_[tree_.dof[i]]*S);
        }
        acc[i]=a;
    }
    for(int i=0;i<n;i++){
        Inertia I=link_inertia(tree_,i);
        Sp v=zeroVel?sp_zero():linkVel_[i];
        f[i]=I*acc[i]+cross_force(v,I*v);
    }
    for(int i=n-1;i>=1;i--){
        if(tree_.dof[i]>=0) out[6+tree_.dof[i]]=dot(link_S(tree_,i),f[i]);
        f[tree_.parent[i]]=f[tree_.parent[i]]+xform_force(pXc_[i],f[i]);
    }
    V3 fl=world_H_base_.R*f[0].lin, fa=world_H_base_.R*f[0].ang;
    for(int k=0;k<3;k++){ out[k]=fl(k); out[3+k]=fa(k); }
}

// KinDynComputations.cpp:2829-2885; mixed "zero" base acceleration -> body (:240-261)
void KinDyn::generalizedBiasForces(double* h){
    Sp a; a.lin=-cross(baseVelBody_.ang,baseVelBody_.lin)-gravityBase_; a.ang=v3(0,0,0);
    rnea(a,nullptr,false,h);
}
// KinDynComputations.cpp:2911-2947
void KinDyn::generalizedGravityForces(double* g){
    Sp a; a.lin=-gravityBase_; a.ang=v3(0,0,0);
    rnea(a,nullptr,true,g);
}
// inverseDynamics in MIXED with zero external forces (KinDynComputations.cpp:1500-1560), used only by
// the identity test  M*nu_dot + h == ID(nu_dot)  (KinDynComputationsUnitTest.cpp:256-279)
void KinDyn::inverseDynamicsMixed(const double* aM,const double* ddq,double* tau){
    M3 Rt=transpose(world_H_base_.R);
    Sp a; a.lin=Rt*v3(aM[0],aM[1],aM[2])-cross(baseVelBody_.ang,baseVelBody_.lin)-gravityBase_; a.ang=Rt*v3(aM[3],aM[4],aM[5]);
    rnea(a,ddq,false,tau);
}

}  // namespace oracle
