// whole_body_controller_ros.hpp -- ROS-facing drop-in for the reference's WholeBodyController node class.
//
// Same class name, same public methods, same topics, parameters and message types as include/anymal_wbc/whole_body_controller.hpp:35-171 /
// src/whole_body_controller.cpp of the reference; everything numerical (iDynTree, the QP assembly, qpOASES) is replaced by ONE robot of the
// batched CUDA library behind the C ABI (wbc_b200.h, N = 1).  A maintainer swaps the include of the node main
// (src/whole_body_controller_node.cpp) for this header and links -lwbc_b200 instead of iDynTree / qpOASES / Eigen.
//
//   reference member (file:line)                                   here
//   WholeBodyController()                     wbc.cpp:22-59        same publishers / subscribers / loadParameters / setInitialState; wbc_create(1, ...)
//   loadParameters()                          wbc.cpp:122-148      same 12 /anymal_wbc/* parameters
//   referenceCallback(WbcReferenceMsg)        wbc.cpp:150-185      unpacks the six arrays + footContacts -> wbc_set_reference (switching derived like :176-184)
//   floatingBaseStateCallback(ModelStates)    wbc.cpp:187-230      model index lookup on the first message, then pose + twist -> wbc_set_state
//   jointStateCallback(JointState)            wbc.cpp:232-254      joint-name permutation on the first message (model order LH,LF,RF,RH) -> wbc_set_state
//   updateState / solveQP / computeJointTorques   wbc.cpp:256-577  wbc_update_state / wbc_solve_qp / wbc_compute_joint_torques + the two publishers of :558-576
//   resetRobotSimState / publishTransform / terminate / controlLoop / run     wbc.cpp:579-683     same loops, same messages
//
// The ROS headers are the real ones in a catkin workspace; with -DWBC_ROS_MOCK the minimal stand-ins of tests/mock_ros are used so that this file
// is compiled (and its control loop run against the GPU) by the test-suite of a machine without ROS.  No CPU fallback: the constructor reports
// through ROS_ERROR_STREAM and leaves initStatus_ false if the CUDA library cannot be used, exactly like a failed model load upstream.
#ifndef ANYMAL_WBC_B200_WHOLE_BODY_CONTROLLER_ROS_HPP
#define ANYMAL_WBC_B200_WHOLE_BODY_CONTROLLER_ROS_HPP

#include <ros/ros.h>
#include <ros/package.h>
#include <std_msgs/Float64MultiArray.h>
#include <geometry_msgs/Pose.h>
#include <gazebo_msgs/ModelStates.h>
#include <gazebo_msgs/ModelState.h>
#include <sensor_msgs/JointState.h>
#include <boost/thread.hpp>
#include <tf/transform_broadcaster.h>
#include "anymal_wbc/WbcReferenceMsg.h"

#include <cmath>
#include <string>
#include <vector>

#include "wbc_b200.h"

const std::string modelName = "anymalModel";                  // wbc.hpp:26
const int numberOfJoints = WBC_NUM_JOINTS;                    // wbc.hpp:27
const int numberOfLegs = WBC_NUM_LEGS;                        // wbc.hpp:28
const double gravityAcceleration = 9.81;                      // wbc.hpp:29
const int qpNumberOfVariables = WBC_QP_NV;                    // wbc.hpp:30
const int qpNumberOfConstraints = WBC_QP_NC;                  // wbc.hpp:31

class WholeBodyController
{
    public:
        WholeBodyController() : initStatus_(1), firstJointStateCallback_(true), firstFloatingBaseStateCallback(true), firstControllerIteration_(true)
        {
            jointTorquePub_ = nh_.advertise<std_msgs::Float64MultiArray>("/anymal/joint_effort_group_controller/command", 10);
            desiredGroundReactionForcesPub_ = nh_.advertise<std_msgs::Float64MultiArray>("/anymal/desired_ground_reaction_forces", 10);
            centerOfMassPub_ = nh_.advertise<geometry_msgs::Pose>("/anymal/com", 10);
            gazeboPub_ = nh_.advertise<gazebo_msgs::ModelState>("/gazebo/set_model_state", 0);

            floatingBaseStateSub_ = nh_.subscribe("/gazebo/model_states", 0, &WholeBodyController::floatingBaseStateCallback, this);
            jointStateSub_ = nh_.subscribe("/anymal/joint_states", 0, &WholeBodyController::jointStateCallback, this);
            plannerReferenceSub_ = nh_.subscribe("/anymal/reference", 0, &WholeBodyController::referenceCallback, this);

            loadParameters();

            // the 13-body ANYmal table compiled from urdf/anymal.urdf is baked into the library: "loading the model" = creating the handle
            wbc_params p; wbc_default_params(&p);
            p.friction = params.friction; p.loopRate = params.loopRate; p.maxTorque = params.maxTorque; p.kpValue = params.kpValue; p.kpValueZ = params.kpValueZ;
            p.kdValue = params.kdValue; p.kiValue = params.kiValue; p.kpSwingValue = params.kpSwingValue; p.kdSwingValue = params.kdSwingValue; p.slackWeight = params.slackWeight;
            for (int i = 0; i < 6 && i < (int)params.refData.size(); ++i) p.initialReferencePose[i] = params.refData[i];
            initStatus_ = (wbc_create(1, 0, &p, nullptr, &handle_) == WBC_OK);
            if (!initStatus_)
            {
                ROS_ERROR_STREAM("WholeBodyController: the CUDA library could not create a controller on device 0 (no sm_100a GPU? there is no CPU fallback)");
            }

            setInitialState();
            for (int i = 0; i < qpNumberOfVariables; ++i) qpSolution_[i] = 0.0;
        }
        ~WholeBodyController() { if (handle_) wbc_destroy(handle_); }
        WholeBodyController(const WholeBodyController&) = delete;
        WholeBodyController& operator=(const WholeBodyController&) = delete;

        void run()
        {
            boost::thread ctrl_loop_t(&WholeBodyController::controlLoop, this);     // starts the control loop thread (wbc.cpp:678-683)
            ros::spin();
            ctrl_loop_t.join();
        }

        void floatingBaseStateCallback(gazebo_msgs::ModelStates modelStateMsg)
        {
            if (firstFloatingBaseStateCallback)
            {
                modelIndex_ = 0;
                bool found = false;
                while (!found && modelIndex_ < (int)modelStateMsg.name.size())
                {
                    if (modelStateMsg.name[modelIndex_] == modelName) found = true; else modelIndex_++;
                }
                firstFloatingBaseStateCallback = false;
            }
            else
            {
                const auto& P = modelStateMsg.pose[modelIndex_]; const auto& T = modelStateMsg.twist[modelIndex_];
                basePos_[0] = P.position.x; basePos_[1] = P.position.y; basePos_[2] = P.position.z;
                baseQuat_[0] = P.orientation.w; baseQuat_[1] = P.orientation.x; baseQuat_[2] = P.orientation.y; baseQuat_[3] = P.orientation.z;
                baseVel_[0] = T.linear.x; baseVel_[1] = T.linear.y; baseVel_[2] = T.linear.z; baseVel_[3] = T.angular.x; baseVel_[4] = T.angular.y; baseVel_[5] = T.angular.z;
                if (handle_) wbc_set_state(handle_, nullptr, nullptr, basePos_, baseQuat_, baseVel_);
            }
        }

        void jointStateCallback(sensor_msgs::JointState jointStateMsg)
        {
            // make sure that the order of the joints from the msg is the same as the model (LH-LF-RF-RH)
            static const char* modelJointNames[numberOfJoints] = {"LH_HAA", "LH_HFE", "LH_KFE", "LF_HAA", "LF_HFE", "LF_KFE", "RF_HAA", "RF_HFE", "RF_KFE", "RH_HAA", "RH_HFE", "RH_KFE"};
            if (firstJointStateCallback_)
            {
                for (int i = 0; i < numberOfJoints; i++)
                {
                    int k = 0;
                    while (jointStateMsg.name[k] != modelJointNames[i]) k++;
                    jointIndex_[i] = k;
                }
                firstJointStateCallback_ = false;
            }
            for (int i = 0; i < numberOfJoints; i++)
            {
                jointPos_[i] = jointStateMsg.position[jointIndex_[i]];
                jointVel_[i] = jointStateMsg.velocity[jointIndex_[i]];
            }
            if (handle_) wbc_set_state(handle_, jointPos_, jointVel_, nullptr, nullptr, nullptr);
        }

        void referenceCallback(anymal_wbc::WbcReferenceMsg refMsg)
        {
            for (int i = 0; i < 6; i++) { desiredPose_[i] = refMsg.desiredComPose.data[i]; desiredCoMVelocity_[i] = refMsg.desiredComVelocity.data[i]; desiredCoMAcceleration_[i] = refMsg.desiredComAcceleration.data[i]; }
            for (int i = 0; i < 3 * numberOfLegs; i++)
            {
                desiredSwingLegsPosition_[i] = refMsg.desiredSwingLegsPosition.data[i];
                desiredSwingLegsVelocity_[i] = refMsg.desiredSwingLegsVelocity.data[i];
                desiredSwingLegsAcceleration_[i] = refMsg.desiredSwingLegsAcceleration.data[i];
            }
            unsigned char c[numberOfLegs];
            isSwitchingFootState_ = false;
            for (int i = 0; i < numberOfLegs; i++)
            {
                if (footContacts_[i] != (int)refMsg.footContacts[i]) isSwitchingFootState_ = true;
                footContacts_[i] = refMsg.footContacts[i];
                c[i] = footContacts_[i] ? 1 : 0;
            }
            const unsigned char sw = isSwitchingFootState_ ? 1 : 0;
            if (handle_) wbc_set_reference(handle_, desiredPose_, desiredCoMVelocity_, desiredCoMAcceleration_, desiredSwingLegsPosition_, desiredSwingLegsVelocity_,
                                           desiredSwingLegsAcceleration_, c, &sw);
        }

        void updateState() { if (handle_) wbc_update_state(handle_); }

        void setInitialState()
        {
            for (int i = 0; i < numberOfLegs; i++) footContacts_[i] = 1;
            const double q0[numberOfJoints] = {0.0, -0.4, 0.8, 0.0, 0.4, -0.8, 0.0, 0.4, -0.8, 0.0, -0.4, 0.8};
            for (int i = 0; i < numberOfJoints; i++) { jointPos_[i] = q0[i]; jointVel_[i] = 0.0; }
            for (int i = 0; i < 6; i++) baseVel_[i] = 0.0;
            basePos_[0] = 0.0; basePos_[1] = 0.0; basePos_[2] = 0.60; baseQuat_[0] = 1.0; baseQuat_[1] = baseQuat_[2] = baseQuat_[3] = 0.0;
            for (int i = 0; i < 6; ++i) { desiredPose_[i] = i < (int)params.refData.size() ? params.refData[i] : 0.0; desiredCoMVelocity_[i] = 0.0; desiredCoMAcceleration_[i] = 0.0; }
            for (int i = 0; i < 3 * numberOfLegs; i++) { desiredSwingLegsAcceleration_[i] = 0.0; desiredSwingLegsVelocity_[i] = 0.0; desiredSwingLegsPosition_[i] = 0.0; }
            centerOfMassPosition_[0] = centerOfMassPosition_[1] = centerOfMassPosition_[2] = 0.0;
            firstControllerIteration_ = true;
            if (handle_) wbc_reset(handle_, 0, 1);              // the library's setInitialState: same values, old T = I, old Jacobians = 0, integral error = 0, next solve = SQProblem::init
        }

        void loadParameters()
        {
            if (!nh_.getParam("/anymal_wbc/modelName", params.modelName)) ROS_ERROR("Failed to get param 'modelName'");
            if (!nh_.getParam("/anymal_wbc/friction", params.friction)) ROS_ERROR("Failed to get param 'friction'");
            if (!nh_.getParam("/anymal_wbc/kpValue", params.kpValue)) ROS_ERROR("Failed to get param 'kpValue'");
            if (!nh_.getParam("/anymal_wbc/kpValueZ", params.kpValueZ)) ROS_ERROR("Failed to get param 'kpValueZ'");
            if (!nh_.getParam("/anymal_wbc/kdValue", params.kdValue)) ROS_ERROR("Failed to get param 'kdValue'");
            if (!nh_.getParam("/anymal_wbc/kiValue", params.kiValue)) ROS_ERROR("Failed to get param 'kiValue'");
            if (!nh_.getParam("/anymal_wbc/kpSwingValue", params.kpSwingValue)) ROS_ERROR("Failed to get param 'kpSwingValue'");
            if (!nh_.getParam("/anymal_wbc/kdSwingValue", params.kdSwingValue)) ROS_ERROR("Failed to get param 'kdSwingValue'");
            if (!nh_.getParam("/anymal_wbc/loopRate", params.loopRate)) ROS_ERROR("Failed to get param 'loopRate'");
            if (!nh_.getParam("/anymal_wbc/maxTorque", params.maxTorque)) ROS_ERROR("Failed to get param 'maxTorque'");
            if (!nh_.getParam("/anymal_wbc/slackWeight", params.slackWeight)) ROS_ERROR("Failed to get param 'slackWeight'");
            if (!nh_.getParam("/anymal_wbc/initialReferencePose", params.refData)) ROS_ERROR("Failed to get param 'initialReferencePose'");
        }

        void computeJointTorques()
        {
            double tau[numberOfJoints] = {0}, grf[3 * numberOfLegs] = {0};
            if (handle_)
            {
                wbc_compute_joint_torques(handle_);
                wbc_get_outputs(handle_, tau, grf, qpSolution_, &qpReturnValue_, nullptr, nullptr, centerOfMassPosition_);     // host pointers: synchronises the stream
            }
            std_msgs::Float64MultiArray desiredgrfMsg;
            for (int i = 0; i < 3 * numberOfLegs; i++) desiredgrfMsg.data.push_back(grf[i]);
            desiredGroundReactionForcesPub_.publish(desiredgrfMsg);
            std_msgs::Float64MultiArray commandMsg;
            // the library zeroes the torques of a robot whose QP did not solve; the reference publishes whatever the stale solution gives and then terminate()s
            // with zeros one line later (wbc.cpp:654-659) -- the message that reaches the actuators last is the same
            for (int i = 0; i < numberOfJoints; i++) commandMsg.data.push_back(tau[i]);
            jointTorquePub_.publish(commandMsg);
        }
        void solveQP() { if (handle_) wbc_solve_qp(handle_); firstControllerIteration_ = false; }

        void resetRobotSimState()
        {
            ros::Rate rosRate(params.loopRate);
            double time = 0.0;
            const double deltaTime = 1.0 / params.loopRate;
            double resetTime = 0.5;
            double zOffset = 0.05;
            while (time <= resetTime)                            // needed to let the controllers load in gazebo
            {
                gazebo_msgs::ModelState stateMsg;
                stateMsg.model_name = params.modelName;
                stateMsg.reference_frame = "world";
                geometry_msgs::Pose pose;
                pose.position.x = desiredPose_[0];
                pose.position.y = desiredPose_[1];
                pose.position.z = desiredPose_[2] + zOffset;
                stateMsg.pose = pose;
                gazeboPub_.publish(stateMsg);
                setInitialState();
                rosRate.sleep();
                time = time + deltaTime;
            }
        }

        void controlLoop()
        {
            ros::Rate rosRate(params.loopRate);
            resetRobotSimState();
            ROS_INFO_STREAM("Loaded reference pose: " << desiredPose_[0] << " " << desiredPose_[1] << " " << desiredPose_[2] << " " << desiredPose_[3] << " " << desiredPose_[4] << " " << desiredPose_[5]);
            double time = 0.0;
            const double deltaTime = 1.0 / params.loopRate;
            long iteration = 0;
            while (ros::ok())
            {
                updateState();
                solveQP();
                computeJointTorques();
                if (qpReturnValue_ != 0)                         // qpOASES::SUCCESSFUL_RETURN
                {
                    ROS_WARN("QP did not solve!");
                    terminate();
                    break;
                }
                geometry_msgs::Pose comMsg;
                comMsg.position.x = centerOfMassPosition_[0];
                comMsg.position.y = centerOfMassPosition_[1];
                comMsg.position.z = centerOfMassPosition_[2];
                centerOfMassPub_.publish(comMsg);
                if (iteration % 10 == 0) publishTransform();
                rosRate.sleep();
                time = time + deltaTime;
            }
        }

        void publishTransform()
        {
            transform_.setOrigin(tf::Vector3(basePos_[0], basePos_[1], basePos_[2]));
            // Eigen::Quaterniond::toRotationMatrix (not normalised) followed by eulAnglesRPY (wbc.cpp:12-20), rows 1 and 2 only
            const double w = baseQuat_[0], x = baseQuat_[1], y = baseQuat_[2], z = baseQuat_[3];
            const double R00 = 1.0 - 2.0 * (y * y + z * z), R10 = 2.0 * (x * y + w * z), R20 = 2.0 * (x * z - w * y), R21 = 2.0 * (y * z + w * x), R22 = 1.0 - 2.0 * (x * x + y * y);
            const double roll = std::atan2(R21, R22), pitch = std::atan2(-R20, std::sqrt(R21 * R21 + R22 * R22)), yaw = std::atan2(R10, R00);
            tf::Quaternion q;
            q.setRPY(roll, pitch, yaw);
            transform_.setRotation(q);
            static ros::Time last_broadcast_time = ros::Time(0);
            ros::Time now = ros::Time::now();
            if (now == ros::Time(0) || now <= last_broadcast_time) now = last_broadcast_time + ros::Duration(1e-6);     // monotonically increasing stamps
            last_broadcast_time = now;
            broadcaster_.sendTransform(tf::StampedTransform(transform_, now, "map", "base"));
        }

        void terminate()
        {
            ROS_INFO("Shutting down whole body controller...");
            std_msgs::Float64MultiArray commandMsg;
            for (int i = 0; i < numberOfJoints; i++) commandMsg.data.push_back(0.0);
            jointTorquePub_.publish(commandMsg);
        }

        bool initStatus() const { return initStatus_; }
        int qpReturnValue() const { return qpReturnValue_; }
        wbc_handle* handle() { return handle_; }

    private:
        ros::NodeHandle nh_;

        ros::Publisher desiredGroundReactionForcesPub_;
        ros::Publisher jointTorquePub_;
        ros::Publisher centerOfMassPub_;
        ros::Publisher gazeboPub_;

        ros::Subscriber floatingBaseStateSub_;
        ros::Subscriber jointStateSub_;
        ros::Subscriber plannerReferenceSub_;

        bool initStatus_;
        bool firstJointStateCallback_;
        bool firstFloatingBaseStateCallback;
        bool firstControllerIteration_;

        struct {                                                  // wbc.hpp:91-104
            std::string modelName;
            double friction = 1.0, loopRate = 400.0, maxTorque = 80.0, kpValue = 6000.0, kpValueZ = 10000.0, kdValue = 1800.0, kiValue = 0.0, kpSwingValue = 250.0,
                   kdSwingValue = 20.0, slackWeight = 1000.0;     // config/params_controller.yaml, in case a parameter is missing
            std::vector<double> refData;
        } params;

        wbc_handle* handle_ = nullptr;                            // replaces mdlLoader_, kinDynComp_, model_, quadraticProblem_ and every Eigen member of wbc.hpp:106-166
        int jointIndex_[numberOfJoints];
        int modelIndex_ = 0;

        double basePos_[3], baseQuat_[4], baseVel_[6], jointPos_[numberOfJoints], jointVel_[numberOfJoints];
        double desiredPose_[6], desiredCoMVelocity_[6], desiredCoMAcceleration_[6];
        double desiredSwingLegsAcceleration_[3 * numberOfLegs], desiredSwingLegsVelocity_[3 * numberOfLegs], desiredSwingLegsPosition_[3 * numberOfLegs];
        double centerOfMassPosition_[3];
        int footContacts_[numberOfLegs];
        bool isSwitchingFootState_ = false;
        double qpSolution_[qpNumberOfVariables];
        int qpReturnValue_ = 0;

        tf::TransformBroadcaster broadcaster_;
        tf::Transform transform_;
};

#endif // ANYMAL_WBC_B200_WHOLE_BODY_CONTROLLER_ROS_HPP
