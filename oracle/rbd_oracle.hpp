// ORACLE -- TEST INFRASTRUCTURE ONLY. Never linked into or called by the product path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg use it.
//
// CPU restatement of the iDynTree 12.3.2 subset that WholeBodyController::updateState() calls
// (iDynTree cannot be built here: it needs Eigen3 and libxml2, neither is in the image).
// Every function cites the reference file:line it follows (paths relative to /root/reference/).
// Parity status: iDynTree ships no golden numbers for ANYmal; this restatement is pinned by the
// identities of lib/idyntree/src/high-level/tests/KinDynComputationsUnitTest.cpp:256-279,506-541
// (tests/test_oracle_rbd.py) and by the URDF known answers (total mass 51.63574).
//
// Conventions (lib/idyntree/src/core): 6-vectors are [linear; angular]; a Transform A_X_B = (R, p)
// maps B coordinates to A coordinates, p = origin of B expressed in A; link twists are body-fixed
// (left-trivialised); the user-facing representation is MIXED (KinDynComputations.cpp:186).
#pragma once
#include <cmath>
#include <cstring>
#include <vector>

namespace oracle {

struct V3 { double x[3]; double& operator()(int i){return x[i];} double operator()(int i) const {return x[i];} };
inline V3 v3(double a,double b,double c){ V3 r{{a,b,c}}; return r; }
inline V3 operator+(const V3&a,const V3&b){ return v3(a(0)+b(0),a(1)+b(1),a(2)+b(2)); }
inline V3 operator-(const V3&a,const V3&b){ return v3(a(0)-b(0),a(1)-b(1),a(2)-b(2)); }
inline V3 operator-(const V3&a){ return v3(-a(0),-a(1),-a(2)); }
inline V3 operator*(double s,const V3&a){ return v3(s*a(0),s*a(1),s*a(2)); }
inline V3 cross(const V3&a,const V3&b){ return v3(a(1)*b(2)-a(2)*b(1), a(2)*b(0)-a(0)*b(2), a(0)*b(1)-a(1)*b(0)); }
inline double dot(const V3&a,const V3&b){ return a(0)*b(0)+a(1)*b(1)+a(2)*b(2); }

struct M3 { double m[9]; double& operator()(int i,int j){return m[3*i+j];} double operator()(int i,int j) const {return m[3*i+j];} };
inline M3 m3_identity(){ M3 r{{1,0,0,0,1,0,0,0,1}}; return r; }
inline M3 m3_zero(){ M3 r{{0,0,0,0,0,0,0,0,0}}; return r; }
inline M3 operator*(const M3&a,const M3&b){ M3 r; for(int i=0;i<3;i++)for(int j=0;j<3;j++){double s=0;for(int k=0;k<3;k++)s+=a(i,k)*b(k,j);r(i,j)=s;} return r; }
inline V3 operator*(const M3&a,const V3&b){ V3 r; for(int i=0;i<3;i++) r(i)=a(i,0)*b(0)+a(i,1)*b(1)+a(i,2)*b(2); return r; }
inline M3 operator+(const M3&a,const M3&b){ M3 r; for(int i=0;i<9;i++) r.m[i]=a.m[i]+b.m[i]; return r; }
inline M3 transpose(const M3&a){ M3 r; for(int i=0;i<3;i++)for(int j=0;j<3;j++) r(i,j)=a(j,i); return r; }
inline M3 skew(const V3&v){ M3 r{{0,-v(2),v(1), v(2),0,-v(0), -v(1),v(0),0}}; return r; }

// core/src/Rotation.cpp RotAxis: Rodrigues formula about a unit axis
inline M3 rot_axis(const V3& u, double th){
    double c=std::cos(th), s=std::sin(th), t=1-c; M3 r;
    r(0,0)=c+u(0)*u(0)*t;      r(0,1)=u(0)*u(1)*t-u(2)*s; r(0,2)=u(0)*u(2)*t+u(1)*s;
    r(1,0)=u(1)*u(0)*t+u(2)*s; r(1,1)=c+u(1)*u(1)*t;      r(1,2)=u(1)*u(2)*t-u(0)*s;
    r(2,0)=u(2)*u(0)*t-u(1)*s; r(2,1)=u(2)*u(1)*t+u(0)*s; r(2,2)=c+u(2)*u(2)*t;
    return r;
}

struct Transform { M3 R; V3 p; };
inline Transform tf_identity(){ Transform t; t.R=m3_identity(); t.p=v3(0,0,0); return t; }
// core/src/Transform.cpp compose / inverse
inline Transform operator*(const Transform&a,const Transform&b){ Transform r; r.R=a.R*b.R; r.p=a.R*b.p+a.p; return r; }
inline Transform inverse(const Transform&a){ Transform r; r.R=transpose(a.R); r.p=-(r.R*a.p); return r; }

struct Sp { V3 lin, ang; };   // spatial motion OR force vector, [linear; angular]
inline Sp sp_zero(){ Sp s; s.lin=v3(0,0,0); s.ang=v3(0,0,0); return s; }
inline Sp operator+(const Sp&a,const Sp&b){ Sp r; r.lin=a.lin+b.lin; r.ang=a.ang+b.ang; return r; }
inline Sp operator-(const Sp&a,const Sp&b){ Sp r; r.lin=a.lin-b.lin; r.ang=a.ang-b.ang; return r; }
inline Sp operator*(double s,const Sp&a){ Sp r; r.lin=s*a.lin; r.ang=s*a.ang; return r; }
inline double dot(const Sp&a,const Sp&b){ return dot(a.lin,b.lin)+dot(a.ang,b.ang); }

// core/src/Transform.cpp transform of a twist / spatial acceleration (motion vector)
inline Sp xform_motion(const Transform&X,const Sp&v){ Sp r; r.ang=X.R*v.ang; r.lin=X.R*v.lin+cross(X.p,r.ang); return r; }
// core/src/Transform.cpp transform of a wrench / momentum (force vector)
inline Sp xform_force(const Transform&X,const Sp&f){ Sp r; r.lin=X.R*f.lin; r.ang=X.R*f.ang+cross(X.p,r.lin); return r; }
// core/src/Twist.cpp Twist*Twist (motion cross product) and Twist*Momentum (force cross product)
inline Sp cross_motion(const Sp&v,const Sp&w){ Sp r; r.lin=cross(v.ang,w.lin)+cross(v.lin,w.ang); r.ang=cross(v.ang,w.ang); return r; }
inline Sp cross_force(const Sp&v,const Sp&f){ Sp r; r.lin=cross(v.ang,f.lin); r.ang=cross(v.lin,f.lin)+cross(v.ang,f.ang); return r; }

// core/src/SpatialInertia.cpp: mass, first moment m*c, rotational inertia about the frame origin
struct Inertia { double m; V3 mc; M3 I; };
inline Inertia inertia_zero(){ Inertia r; r.m=0; r.mc=v3(0,0,0); r.I=m3_zero(); return r; }
inline Inertia operator+(const Inertia&a,const Inertia&b){ Inertia r; r.m=a.m+b.m; r.mc=a.mc+b.mc; r.I=a.I+b.I; return r; }
// SpatialInertia.cpp operator*(Twist): momentum = [m v - mc x w ; mc x v + I w]
inline Sp operator*(const Inertia&I,const Sp&v){ Sp r; r.lin=I.m*v.lin-cross(I.mc,v.ang); r.ang=cross(I.mc,v.lin)+I.I*v.ang; return r; }
// Transform.cpp operator*(SpatialInertia): rotate, then shift the origin by p
inline Inertia xform_inertia(const Transform&X,const Inertia&B){
    Inertia A; A.m=B.m; V3 Rmc=X.R*B.mc; A.mc=Rmc+B.m*X.p;
    M3 RI=X.R*B.I*transpose(X.R);
    // I_A = R I_B R^T - S(p) S(R mc) - S(R mc + m p) S(p)
    M3 Sp_=skew(X.p), Sc=skew(Rmc), Sa=skew(A.mc);
    M3 t1=Sp_*Sc, t2=Sa*Sp_;
    for(int i=0;i<9;i++) A.I.m[i]=RI.m[i]-t1.m[i]-t2.m[i];
    return A;
}

struct Tree {
    int n=0; const int *parent=nullptr,*dof=nullptr;
    const double *restR=nullptr,*restP=nullptr,*axis=nullptr,*mass=nullptr,*mcom=nullptr,*inertia=nullptr;
    const int* foot_link=nullptr; const double *footR=nullptr,*footP=nullptr;
    double total_mass=0;
};
const Tree& anymal_full_tree();    // 78 links (rbd_oracle.cpp includes oracle/anymal_tree_full.h)
const Tree& anymal_lumped_tree();  // 13 bodies (csrc/anymal_model.h) -- same table the CUDA kernels bake in

// One KinDynComputations-like evaluator (MIXED representation only).
class KinDyn {
public:
    explicit KinDyn(const Tree& t);
    // KinDynComputations.cpp:752-839
    void setRobotState(const Transform& world_H_base, const double* q, const double* baseVelMixed, const double* dq, const V3& gravity);
    V3 getCenterOfMassPosition();                               // :1915-1924
    V3 getCenterOfMassVelocity();                               // :1941-1958
    void getFreeFloatingMassMatrix(double* M18x18_rowmajor);    // :2606-2630
    void getFrameFreeFloatingJacobian(int foot, double* J6x18_rowmajor);   // :1607-1679 + Jacobians.cpp:16-52
    Transform getWorldTransform(int foot);                      // :1299-1340
    Sp getFrameVel(int foot);                                   // :1409-1445
    void generalizedBiasForces(double* h18);                    // :2829-2885
    void generalizedGravityForces(double* g18);                 // :2911-2947
    double totalMass() const { return tree_.total_mass; }
    // oracle-side instrumentation for the iDynTree-style identities
    void inverseDynamicsMixed(const double* baseAccMixed6, const double* ddq, double* tau18);
private:
    void computeFwdKinematics();                                // :372-388 -> ForwardKinematics.cpp:144-187
    void computeRawMassMatrixAndTotalMomentum();                // :390-418 -> Dynamics.cpp:146-273, :26-42
    Transform jointTransform(int link) const;                   // parent_X_child, RevoluteJoint.cpp:128-134 / FixedJoint
    void rnea(const Sp& baseProperAcc, const double* ddq, bool zeroVel, double* out18);
    const Tree& tree_;
    Transform world_H_base_; std::vector<double> q_, dq_; Sp baseVelBody_; V3 gravityBase_;
    bool fkOk_=false, massOk_=false;
    std::vector<Transform> linkPos_, pXc_; std::vector<Sp> linkVel_; std::vector<Inertia> crb_;
    double rawM_[18*18]; Sp totalMomentum_;
};

}  // namespace oracle
