// ORACLE -- TEST INFRASTRUCTURE ONLY (see wbc_oracle.hpp). Restates src/whole_body_controller.cpp of the reference.
#include "wbc_oracle.hpp"
#include <qpOASES.hpp>

namespace oracle {

// wbc.cpp:12-20
static void eulAnglesRPY(const M3& R,double* rpy){
    rpy[0]=std::atan2(R(2,1),R(2,2));
    rpy[1]=std::atan2(-R(2,0),std::sqrt(R(2,1)*R(2,1)+R(2,2)*R(2,2)));
    rpy[2]=std::atan2(R(1,0),R(0,0));
}

WbcOracle::WbcOracle(const Tree& tree,const Params& p): params(p), kinDyn_(tree) {
    quadraticProblem_.reset(new qpOASES::SQProblem(qpNumberOfVariables,qpNumberOfConstraints));   // wbc.cpp:24
    setInitialState();
    for(int i=0;i<42;i++) qpSolution_[i]=0;                                                      // wbc.cpp:58
    for(int i=0;i<112;i++){ dual_[i]=0; workingSet_[i]=0; }
    for(int i=0;i<12;i++){ tau[i]=0; grf[i]=0; }
}
WbcOracle::~WbcOracle(){}

// wbc.cpp:65-120. transformationMatrixDotInverse_ is never initialised by the reference; defined as 0 here.
void WbcOracle::setInitialState(){
    for(int i=0;i<4;i++) footContacts_[i]=true;
    totalMass_=kinDyn_.totalMass();
    T_world_base_=tf_identity();
    const double q0[12]={0.0,-0.4,0.8, 0.0,0.4,-0.8, 0.0,0.4,-0.8, 0.0,-0.4,0.8};
    for(int i=0;i<12;i++){ jointPos_[i]=q0[i]; jointVel_[i]=0; }
    for(int i=0;i<6;i++) baseVel_[i]=0;
    T_world_base_.p=v3(0,0,0.60);
    massMatrix_=Mat<18,18>::identity(); transformationMatrix_=Mat<18,18>::identity(); oldTransformationMatrix_=Mat<18,18>::identity();
    oldCentroidStanceJacobian_=Mat<12,18>(); oldCentroidSwingJacobian_=Mat<12,18>();
    transformationMatrixDot_=Mat<18,18>(); transformationMatrixDotInverse_=Mat<18,18>();
    centroidStanceJacobianDot_=Mat<12,18>(); centroidSwingJacobianDot_=Mat<12,18>();
    for(int i=0;i<3;i++) centerOfMassPosition_[i]=0;
    for(int i=0;i<6;i++){ desiredPose_[i]=params.refData[i]; desiredCoMVelocity_[i]=0; desiredCoMAcceleration_[i]=0; integralError_[i]=0; }
    for(int i=0;i<12;i++){ desiredSwingLegsAcceleration_[i]=0; desiredSwingLegsVelocity_[i]=0; desiredSwingLegsPosition_[i]=0; }
    isSwitchingFootState_=false;
}

void WbcOracle::setJointState(const double* q,const double* dq){ for(int i=0;i<12;i++){ jointPos_[i]=q[i]; jointVel_[i]=dq[i]; } }

// wbc.cpp:209-228; Eigen::Quaterniond(w,x,y,z).toRotationMatrix() (no normalisation)
void WbcOracle::setFloatingBaseState(const double* pos,const double* qwxyz,const double* tw){
    double w=qwxyz[0],x=qwxyz[1],y=qwxyz[2],z=qwxyz[3];
    double tx=2*x,ty=2*y,tz=2*z, twx=tx*w,twy=ty*w,twz=tz*w, txx=tx*x,txy=ty*x,txz=tz*x, tyy=ty*y,tyz=tz*y,tzz=tz*z;
    M3& R=T_world_base_.R;
    R(0,0)=1-(tyy+tzz); R(0,1)=txy-twz; R(0,2)=txz+twy;
    R(1,0)=txy+twz; R(1,1)=1-(txx+tzz); R(1,2)=tyz-twx;
    R(2,0)=txz-twy; R(2,1)=tyz+twx; R(2,2)=1-(txx+tyy);
    T_world_base_.p=v3(pos[0],pos[1],pos[2]);
    for(int i=0;i<6;i++) baseVel_[i]=tw[i];          // [linear world ; angular world] (:228)
}

// wbc.cpp:150-185
void WbcOracle::setReference(const double* pose,const double* vel,const double* acc,const double* sp,const double* sv,const double* sa,
                             const unsigned char* c,int switching){
    for(int i=0;i<6;i++){ desiredPose_[i]=pose[i]; desiredCoMVelocity_[i]=vel[i]; desiredCoMAcceleration_[i]=acc[i]; }
    for(int i=0;i<12;i++){ desiredSwingLegsPosition_[i]=sp[i]; desiredSwingLegsVelocity_[i]=sv[i]; desiredSwingLegsAcceleration_[i]=sa[i]; }
    bool sw=false;
    for(int i=0;i<4;i++){ if(footContacts_[i]!=(c[i]!=0)) sw=true; footContacts_[i]=(c[i]!=0); }
    isSwitchingFootState_ = (switching<0)? sw : (switching!=0);
}

// wbc.cpp:296-320
Mat<18,18> WbcOracle::computeTransformationMatrix(){
    V3 d=v3(centerOfMassPosition_[0]-T_world_base_.p(0),centerOfMassPosition_[1]-T_world_base_.p(1),centerOfMassPosition_[2]-T_world_base_.p(2));
    Mat<6,6> Xinv=Mat<6,6>::identity(); M3 S=skew(d);
    for(int i=0;i<3;i++)for(int j=0;j<3;j++) Xinv(i,3+j)=-S(i,j);
    Mat<6,18> sel; for(int i=0;i<6;i++) sel(i,i)=1;
    Mat<6,18> JG=Xinv*inverse(massMatrixBase_)*sel*massMatrix_;
    Mat<18,18> T; for(int i=0;i<6;i++)for(int j=0;j<18;j++) T(i,j)=JG(i,j);
    for(int i=0;i<12;i++) T(6+i,6+i)=1;
    return T;
}

// wbc.cpp:322-342 (leg order LH, LF, RF, RH; rows 0..2 of each 6x18 MIXED jacobian)
void WbcOracle::computeJacobians(Mat<12,18>& st,Mat<12,18>& sw){
    double J[6*18];
    for(int l=0;l<4;l++){
        kinDyn_.getFrameFreeFloatingJacobian(l,J);
        for(int r=0;r<3;r++)for(int c=0;c<18;c++){
            footJacobianLin_(3*l+r,c)=J[r*18+c];
            st(3*l+r,c)=J[r*18+c]*(footContacts_[l]?1.0:0.0);
            sw(3*l+r,c)=J[r*18+c]*(footContacts_[l]?0.0:1.0);
        }
    }
}

// wbc.cpp:544-551
Mat<18,1> WbcOracle::computeCoriolisBias(){
    double h[18],gg[18]; kinDyn_.generalizedBiasForces(h); kinDyn_.generalizedGravityForces(gg);
    Mat<18,1> r; for(int i=0;i<18;i++) r(i,0)=h[i]-gg[i]; return r;
}

// wbc.cpp:384-402
void WbcOracle::computeDerivatives(){
    if(isSwitchingFootState_){
        transformationMatrixDot_=Mat<18,18>(); centroidStanceJacobianDot_=Mat<12,18>(); centroidSwingJacobianDot_=Mat<12,18>();
    } else {
        const double timeStep=1.0/params.loopRate;
        for(int i=0;i<18*18;i++) transformationMatrixDot_.a[i]=(transformationMatrix_.a[i]-oldTransformationMatrix_.a[i])/timeStep;
        for(int i=0;i<12*18;i++){ centroidStanceJacobianDot_.a[i]=(centroidStanceJacobian_.a[i]-oldCentroidStanceJacobian_.a[i])/timeStep;
                                  centroidSwingJacobianDot_.a[i]=(centroidSwingJacobian_.a[i]-oldCentroidSwingJacobian_.a[i])/timeStep; }
    }
    oldTransformationMatrix_=transformationMatrix_; oldCentroidStanceJacobian_=centroidStanceJacobian_; oldCentroidSwingJacobian_=centroidSwingJacobian_;
}

// wbc.cpp:256-294
void WbcOracle::updateState(){
    kinDyn_.setRobotState(T_world_base_,jointPos_,baseVel_,jointVel_,v3(0,0,gravityAcceleration));        // :258, gravity_ :78
    V3 com=kinDyn_.getCenterOfMassPosition(), vcom=kinDyn_.getCenterOfMassVelocity();
    for(int i=0;i<3;i++){ centerOfMassPosition_[i]=com(i); centerOfMassVelocity_[i]=vcom(i); centerOfMassVelocity_[3+i]=baseVel_[3+i]; }
    double rpy[3]; eulAnglesRPY(T_world_base_.R,rpy);
    for(int i=0;i<3;i++){ currentPose_[i]=com(i); currentPose_[3+i]=rpy[i]; }
    kinDyn_.getFreeFloatingMassMatrix(massMatrix_.a.data());
    massMatrixBase_=block<18,18,6,6>(massMatrix_,0,0);
    transformationMatrix_=computeTransformationMatrix();
    centroidMassMatrix_=transpose(inverse(transformationMatrix_))*massMatrix_*inverse(transformationMatrix_);     // :270
    centroidMassMatrixBase_=block<18,18,6,6>(centroidMassMatrix_,0,0);
    centroidMassMatrixJoints_=block<18,18,12,12>(centroidMassMatrix_,6,6);
    Mat<12,18> stanceJ,swingJ; computeJacobians(stanceJ,swingJ);
    centroidStanceJacobian_=stanceJ*inverse(transformationMatrix_);                                               // :278
    centroidSwingJacobian_=swingJ*inverse(transformationMatrix_);                                                 // :282
    Mat<18,1> nu; for(int i=0;i<6;i++) nu(i,0)=baseVel_[i]; for(int i=0;i<12;i++) nu(6+i,0)=jointVel_[i];
    coriolisBias_=computeCoriolisBias();
    centroidGeneralizedBias_=transpose(inverse(transformationMatrix_))*(coriolisBias_+massMatrix_*transformationMatrixDotInverse_*nu);   // :289
    computeDerivatives();
    transformationMatrixDotInverse_=(-1.0)*(inverse(transformationMatrix_)*transformationMatrixDot_*inverse(transformationMatrix_));    // :293
}

// wbc.cpp:426-445 (side effect: integral error)
void WbcOracle::computeDesiredWrench(double* W){
    for(int i=0;i<6;i++){
        double kp=(i==2)?params.kpValueZ:params.kpValue;
        double mcb=0; for(int j=0;j<6;j++) mcb+=centroidMassMatrixBase_(i,j)*desiredCoMAcceleration_[j];
        double gw=(i==2)?totalMass_*gravityAcceleration:0.0;
        W[i]= -kp*(currentPose_[i]-desiredPose_[i]) - params.kdValue*(centerOfMassVelocity_[i]-desiredCoMVelocity_[i])
              - params.kiValue*integralError_[i] + gw + mcb;
    }
    for(int i=0;i<6;i++) integralError_[i]=integralError_[i]+(currentPose_[i]-desiredPose_[i])/params.loopRate;
}

// wbc.cpp:447-464 with computeSwingFootPosition/Velocity (:344-382)
void WbcOracle::computeCommandedAccelerationSwingLegs(double* r){
    for(int l=0;l<4;l++){
        Transform H=kinDyn_.getWorldTransform(l); Sp v=kinDyn_.getFrameVel(l);
        for(int k=0;k<3;k++){ swingFootPosition_[3*l+k]=H.p(k); swingFootVelocity_[3*l+k]=v.lin(k); }
    }
    for(int i=0;i<12;i++)
        r[i]=desiredSwingLegsAcceleration_[i]+params.kdSwingValue*(desiredSwingLegsVelocity_[i]-swingFootVelocity_[i])
             +params.kpSwingValue*(desiredSwingLegsPosition_[i]-swingFootPosition_[i]);
    for(int l=0;l<4;l++) for(int k=0;k<3;k++) r[3*l+k]*=(footContacts_[l]?0.0:1.0);
}

// wbc.cpp:468-515
void WbcOracle::assembleQP(){
    Mat<12,6> Jstb=block<12,18,12,6>(centroidStanceJacobian_,0,0), Jswb=block<12,18,12,6>(centroidSwingJacobian_,0,0);
    Mat<12,12> Jstj=block<12,18,12,12>(centroidStanceJacobian_,0,6), Jswj=block<12,18,12,12>(centroidSwingJacobian_,0,6);
    // H = S^T Jst_b Q Jst_b^T S + R, Q = I6, R = diag(1,...,1, slackWeight x12)                     (:468-482)
    for(int i=0;i<42*42;i++) H[i]=0;
    Mat<12,12> JJt=Jstb*transpose(Jstb);
    for(int i=0;i<12;i++)for(int j=0;j<12;j++) H[(18+i)*42+18+j]=JJt(i,j);
    for(int i=0;i<42;i++) H[i*42+i]+=(i>=30)?params.slackWeight:1.0;
    computeDesiredWrench(desiredWrench_);                                                             // (:483-484)
    for(int i=0;i<42;i++) g[i]=0;
    for(int i=0;i<12;i++){ double s=0; for(int j=0;j<6;j++) s+=Jstb(i,j)*desiredWrench_[j]; g[18+i]=-s; }
    // A (:486-497)
    for(int i=0;i<70*42;i++) A[i]=0;
    const double mu=params.friction;
    const double D[4][3]={{1,0,-mu},{-1,0,-mu},{0,1,-mu},{0,-1,-mu}};                                 // (:404-424)
    for(int i=0;i<6;i++){ for(int j=0;j<6;j++) A[i*42+j]=centroidMassMatrixBase_(i,j); for(int j=0;j<12;j++) A[i*42+18+j]=-Jstb(j,i); }
    for(int i=0;i<12;i++){ for(int j=0;j<6;j++) A[(6+i)*42+j]=Jstb(i,j); for(int j=0;j<12;j++) A[(6+i)*42+6+j]=Jstj(i,j); }
    for(int l=0;l<4;l++)for(int r=0;r<4;r++)for(int c=0;c<3;c++) A[(18+4*l+r)*42+18+3*l+c]=D[r][c]*(footContacts_[l]?1.0:0.0);
    for(int i=0;i<12;i++){ for(int j=0;j<12;j++){ A[(34+i)*42+6+j]=centroidMassMatrixJoints_(i,j); A[(34+i)*42+18+j]=-Jstj(j,i); } }
    for(int i=0;i<12;i++){ for(int j=0;j<6;j++){ A[(46+i)*42+j]=Jswb(i,j); A[(58+i)*42+j]=Jswb(i,j); }
                           for(int j=0;j<12;j++){ A[(46+i)*42+6+j]=Jswj(i,j); A[(58+i)*42+6+j]=Jswj(i,j); }
                           A[(46+i)*42+30+i]=-1.0; A[(58+i)*42+30+i]=1.0; }
    // bounds (:499-515)
    double r[12]; computeCommandedAccelerationSwingLegs(r);
    const double INF=qpOASES::INFTY;
    for(int i=0;i<6;i++){ double gw=(i==2)?totalMass_*gravityAcceleration:0.0; ubA[i]=-gw; lbA[i]=-gw; }
    for(int i=0;i<12;i++){
        double a=0,b=0,c=0,d=0;
        for(int j=0;j<6;j++){ a+=centroidStanceJacobianDot_(i,j)*centerOfMassVelocity_[j]; c+=centroidSwingJacobianDot_(i,j)*centerOfMassVelocity_[j]; }
        for(int j=0;j<12;j++){ b+=centroidStanceJacobianDot_(i,6+j)*jointVel_[j]; d+=centroidSwingJacobianDot_(i,6+j)*jointVel_[j]; }
        ubA[6+i]=-a-b; lbA[6+i]=-a-b;
        ubA[34+i]=params.maxTorque-centroidGeneralizedBias_(6+i,0); lbA[34+i]=-params.maxTorque-centroidGeneralizedBias_(6+i,0);
        ubA[46+i]=r[i]-c-d; lbA[46+i]=-INF;
        ubA[58+i]=INF;      lbA[58+i]=r[i]-c-d;
    }
    for(int i=0;i<16;i++){ ubA[18+i]=0; lbA[18+i]=-INF; }
}

// wbc.cpp:517-541
void WbcOracle::solveQP(){
    assembleQP();
    if(skipSolve) return;
    qpOASES::int_t nWSR=100;
    qpOASES::Options myOptions; myOptions.setToReliable(); myOptions.printLevel=qpOASES::PL_LOW;
    quadraticProblem_->setOptions(myOptions);
    if(firstControllerIteration_){
        qpReturnValue_=quadraticProblem_->init(H,g,A,nullptr,nullptr,lbA,ubA,nWSR,0);
        firstControllerIteration_=false;
    } else {
        qpReturnValue_=quadraticProblem_->hotstart(H,g,A,nullptr,nullptr,lbA,ubA,nWSR,0);
    }
    nWSR_=(int)nWSR;
    qpOASES::real_t xOpt[qpNumberOfVariables];
    // the reference leaves xOpt uninitialised when the QP is not solved (QProblemB.cpp:774-793); keep the previous solution instead
    for(int i=0;i<42;i++) xOpt[i]=qpSolution_[i];
    quadraticProblem_->getPrimalSolution(xOpt);
    for(int i=0;i<42;i++) qpSolution_[i]=xOpt[i];
    quadraticProblem_->getDualSolution(dual_);
    quadraticProblem_->getWorkingSet(workingSet_);
}

// wbc.cpp:553-577
void WbcOracle::computeJointTorques(){
    for(int i=0;i<12;i++) grf[i]=qpSolution_[18+i];
    for(int i=0;i<12;i++){
        double s=0; for(int j=0;j<12;j++) s+=centroidMassMatrixJoints_(i,j)*qpSolution_[6+j];
        double t=0; for(int j=0;j<12;j++) t+=centroidStanceJacobian_(j,6+i)*qpSolution_[18+j];
        tau[i]=s+centroidGeneralizedBias_(6+i,0)-t;
    }
}

}  // namespace oracle
