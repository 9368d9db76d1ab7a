// ORACLE -- TEST INFRASTRUCTURE ONLY. Never linked into or called by the product path.
// Literal CPU restatement of WholeBodyController's numerics (src/whole_body_controller.cpp:3-20,65-120,
// 256-577 and include/anymal_wbc/whole_body_controller.hpp:27-32,91-166 of the reference) on top of
// rbd_oracle (restated iDynTree) and the REAL, unmodified vendored qpOASES 3.2.2 (oracle/_ref/libqpoases_ref.a).
// The reference controller cannot be compiled here (needs ROS, Gazebo msgs, Eigen) and ships no test or
// fixture for this path, so at controller level: PARITY UNPINNED by the reference; the qpOASES half is the
// real library. Dense algebra keeps the reference's operation order, incl. the seven 18x18 inversions.
#pragma once
#include <array>
#include <memory>
#include "rbd_oracle.hpp"

namespace qpOASES { class SQProblem; }

namespace oracle {

template<int R,int C> struct Mat {
    std::array<double,R*C> a{};
    double& operator()(int i,int j){ return a[i*C+j]; }
    double operator()(int i,int j) const { return a[i*C+j]; }
    static Mat identity(){ Mat m; for(int i=0;i<(R<C?R:C);i++) m(i,i)=1; return m; }
};
template<int R,int K,int C> Mat<R,C> operator*(const Mat<R,K>&A,const Mat<K,C>&B){
    Mat<R,C> r; for(int i=0;i<R;i++)for(int k=0;k<K;k++){ double aik=A(i,k); if(aik==0.0) continue; for(int j=0;j<C;j++) r(i,j)+=aik*B(k,j);} return r; }
template<int R,int C> Mat<R,C> operator+(const Mat<R,C>&A,const Mat<R,C>&B){ Mat<R,C> r; for(int i=0;i<R*C;i++) r.a[i]=A.a[i]+B.a[i]; return r; }
template<int R,int C> Mat<R,C> operator-(const Mat<R,C>&A,const Mat<R,C>&B){ Mat<R,C> r; for(int i=0;i<R*C;i++) r.a[i]=A.a[i]-B.a[i]; return r; }
template<int R,int C> Mat<R,C> operator*(double s,const Mat<R,C>&A){ Mat<R,C> r; for(int i=0;i<R*C;i++) r.a[i]=s*A.a[i]; return r; }
template<int R,int C> Mat<C,R> transpose(const Mat<R,C>&A){ Mat<C,R> r; for(int i=0;i<R;i++)for(int j=0;j<C;j++) r(j,i)=A(i,j); return r; }
template<int R,int C,int R2,int C2> Mat<R2,C2> block(const Mat<R,C>&A,int r0,int c0){ Mat<R2,C2> r; for(int i=0;i<R2;i++)for(int j=0;j<C2;j++) r(i,j)=A(r0+i,c0+j); return r; }
// partial-pivoting LU inverse (what Eigen's fixed-size .inverse() does for N > 4: PartialPivLU)
template<int N> Mat<N,N> inverse(const Mat<N,N>&A){
    Mat<N,N> lu=A, inv=Mat<N,N>::identity(); int piv[N];
    for(int k=0;k<N;k++){
        int p=k; double best=std::fabs(lu(k,k));
        for(int i=k+1;i<N;i++) if(std::fabs(lu(i,k))>best){ best=std::fabs(lu(i,k)); p=i; }
        piv[k]=p;
        if(p!=k) for(int j=0;j<N;j++){ std::swap(lu(k,j),lu(p,j)); std::swap(inv(k,j),inv(p,j)); }
        for(int i=k+1;i<N;i++){ double l=lu(i,k)/lu(k,k); lu(i,k)=l; if(l==0.0) continue;
            for(int j=k+1;j<N;j++) lu(i,j)-=l*lu(k,j);
            for(int j=0;j<N;j++) inv(i,j)-=l*inv(k,j); }
    }
    for(int k=N-1;k>=0;k--){ for(int j=0;j<N;j++){ double s=inv(k,j); for(int i=k+1;i<N;i++) s-=lu(k,i)*inv(i,j); inv(k,j)=s/lu(k,k);} }
    (void)piv; return inv;
}

// whole_body_controller.hpp:91-104 + config/params_controller.yaml
struct Params {
    double friction=1.0, loopRate=400.0, maxTorque=80.0, kpValue=6000.0, kpValueZ=10000.0, kdValue=1800.0, kiValue=0.0,
           kpSwingValue=250.0, kdSwingValue=20.0, slackWeight=1000.0;
    double refData[6]={0,0,0.5,0,0,0};
};

constexpr int numberOfJoints=12, numberOfLegs=4, qpNumberOfVariables=42, qpNumberOfConstraints=70;   // wbc.hpp:27-32
constexpr double gravityAcceleration=9.81;

class WbcOracle {
public:
    WbcOracle(const Tree& tree,const Params& p);
    ~WbcOracle();
    void setInitialState();                                      // wbc.cpp:65-120
    // the three ROS callbacks, as plain setters (wbc.cpp:150-254)
    void setJointState(const double* q,const double* dq);        // :249-253 (already in model DOF order)
    void setFloatingBaseState(const double* pos,const double* quat_wxyz,const double* twistLinAng);   // :209-228
    // switching < 0 => derive from the previous contacts like :176-184
    void setReference(const double* pose6,const double* vel6,const double* acc6,const double* swPos12,const double* swVel12,
                      const double* swAcc12,const unsigned char* contacts4,int switching);
    void updateState();                                          // :256-294
    void solveQP();                                              // :466-542
    void computeJointTorques();                                  // :553-577
    void step(){ updateState(); solveQP(); computeJointTorques(); }
    void assembleQP();                                           // the part of solveQP before the qpOASES call (:468-515)
    bool skipSolve=false;                                        // tests: assemble only

    Params params;
    // ---- state (names follow wbc.hpp:114-166) ----
    bool footContacts_[4]; bool isSwitchingFootState_=false; bool firstControllerIteration_=true;
    Transform T_world_base_; double jointPos_[12], jointVel_[12], baseVel_[6];
    double desiredPose_[6], desiredCoMVelocity_[6], desiredCoMAcceleration_[6];
    double desiredSwingLegsPosition_[12], desiredSwingLegsVelocity_[12], desiredSwingLegsAcceleration_[12];
    double totalMass_=0, centerOfMassPosition_[3], centerOfMassVelocity_[6], currentPose_[6], integralError_[6];
    Mat<18,18> massMatrix_, transformationMatrix_, oldTransformationMatrix_, transformationMatrixDot_, transformationMatrixDotInverse_, centroidMassMatrix_;
    Mat<6,6> massMatrixBase_, centroidMassMatrixBase_; Mat<12,12> centroidMassMatrixJoints_;
    Mat<12,18> centroidStanceJacobian_, centroidSwingJacobian_, oldCentroidStanceJacobian_, oldCentroidSwingJacobian_,
               centroidStanceJacobianDot_, centroidSwingJacobianDot_, footJacobianLin_;
    Mat<18,1> centroidGeneralizedBias_, coriolisBias_;
    double swingFootPosition_[12], swingFootVelocity_[12], desiredWrench_[6];
    // ---- QP data (row-major like wbc.cpp:479,486) and results ----
    double H[42*42], g[42], A[70*42], lbA[70], ubA[70];
    double qpSolution_[42], dual_[112], workingSet_[112]; int qpReturnValue_=0, nWSR_=0;
    double tau[12], grf[12];
private:
    KinDyn kinDyn_;
    std::unique_ptr<qpOASES::SQProblem> quadraticProblem_;
    Mat<18,18> computeTransformationMatrix();                    // :296-320
    void computeJacobians(Mat<12,18>& st,Mat<12,18>& sw);         // :322-342
    void computeDerivatives();                                   // :384-402
    Mat<18,1> computeCoriolisBias();                             // :544-551
    void computeDesiredWrench(double* W);                        // :426-445
    void computeCommandedAccelerationSwingLegs(double* r);       // :447-464
};

}  // namespace oracle
