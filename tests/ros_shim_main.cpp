// Test driver of include/anymal_wbc_b200/whole_body_controller_ros.hpp against the mock roscpp of tests/mock_ros (tests/test_ros_shim.py).
// argv: file with 12 q, 12 qd, 3 pos, 4 quat (w x y z), 6 twist, then K = number of control-loop iterations.  The joint-state message lists the joints in
// Gazebo's alphabetical order (LF, LH, RF, RH), so the name permutation of jointStateCallback (wbc.cpp:232-254) is exercised.
// Prints one line per published topic of the LAST iteration.
#include <cstdio>
#include <fstream>
#include "anymal_wbc_b200/whole_body_controller_ros.hpp"
int main(int argc, char** argv) {
    if (argc < 3) return 2;
    std::ifstream f(argv[1]); double q[12], qd[12], pos[3], quat[4], tw[6];
    for (double& v : q) f >> v;
    for (double& v : qd) f >> v;
    for (double& v : pos) f >> v;
    for (double& v : quat) f >> v;
    for (double& v : tw) f >> v;
    const int K = atoi(argv[2]);
    mock::setParam<std::string>("/anymal_wbc/modelName", "anymalModel");
    const char* names[] = {"friction", "kpValue", "kpValueZ", "kdValue", "kiValue", "kpSwingValue", "kdSwingValue", "loopRate", "maxTorque", "slackWeight"};
    const double vals[] = {1.0, 6000.0, 10000.0, 1800.0, 0.0, 250.0, 20.0, 400.0, 80.0, 1000.0};                         // config/params_controller.yaml
    for (int i = 0; i < 10; i++) mock::setParam<double>(std::string("/anymal_wbc/") + names[i], vals[i]);
    mock::setParam<std::vector<double>>("/anymal_wbc/initialReferencePose", {0.0, 0.0, 0.5, 0.0, 0.0, 0.0});
    ros::init(argc, argv, "whole_body_controller_node");
    WholeBodyController wbc;
    if (!wbc.initStatus()) { printf("INIT_FAILED\n"); return 3; }
    // the reset phase of controlLoop() calls setInitialState() 201 times; the state messages arrive afterwards, so the loop is run in two parts:
    mock::Bus::get().okBudget = 0; wbc.controlLoop();                                   // resetRobotSimState + zero iterations
    sensor_msgs::JointState js; const char* order[4] = {"LF", "LH", "RF", "RH"}; const int modelLeg[4] = {1, 0, 2, 3}; const char* jn[3] = {"HAA", "HFE", "KFE"};
    for (int l = 0; l < 4; l++) for (int j = 0; j < 3; j++) { js.name.push_back(std::string(order[l]) + "_" + jn[j]); js.position.push_back(q[3 * modelLeg[l] + j]); js.velocity.push_back(qd[3 * modelLeg[l] + j]); }
    mock::deliver<sensor_msgs::JointState>("/anymal/joint_states", js);
    gazebo_msgs::ModelStates ms; ms.name = {"ground_plane", "anymalModel"}; ms.pose.resize(2); ms.twist.resize(2);
    ms.pose[1].position.x = pos[0]; ms.pose[1].position.y = pos[1]; ms.pose[1].position.z = pos[2];
    ms.pose[1].orientation.w = quat[0]; ms.pose[1].orientation.x = quat[1]; ms.pose[1].orientation.y = quat[2]; ms.pose[1].orientation.z = quat[3];
    ms.twist[1].linear.x = tw[0]; ms.twist[1].linear.y = tw[1]; ms.twist[1].linear.z = tw[2]; ms.twist[1].angular.x = tw[3]; ms.twist[1].angular.y = tw[4]; ms.twist[1].angular.z = tw[5];
    mock::deliver<gazebo_msgs::ModelStates>("/gazebo/model_states", ms);               // first message: model index lookup only (wbc.cpp:189-208)
    mock::deliver<gazebo_msgs::ModelStates>("/gazebo/model_states", ms);
    anymal_wbc::WbcReferenceMsg ref; ref.desiredComPose.data = {0.0, 0.0, 0.5, 0.0, 0.0, 0.0}; ref.desiredComVelocity.data.assign(6, 0.0); ref.desiredComAcceleration.data.assign(6, 0.0);
    ref.desiredSwingLegsPosition.data.assign(12, 0.0); ref.desiredSwingLegsVelocity.data.assign(12, 0.0); ref.desiredSwingLegsAcceleration.data.assign(12, 0.0);
    mock::deliver<anymal_wbc::WbcReferenceMsg>("/anymal/reference", ref);
    for (int k = 0; k < K; k++) { wbc.updateState(); wbc.solveQP(); wbc.computeJointTorques(); if (wbc.qpReturnValue() != 0) { wbc.terminate(); break; } }
    wbc.publishTransform();
    const auto& tau = mock::lastPublished<std_msgs::Float64MultiArray>("/anymal/joint_effort_group_controller/command");
    const auto& grf = mock::lastPublished<std_msgs::Float64MultiArray>("/anymal/desired_ground_reaction_forces");
    printf("rc %d published %d\n", wbc.qpReturnValue(), mock::Bus::get().count["/anymal/joint_effort_group_controller/command"]);
    printf("tau"); for (double v : tau.data) printf(" %.17g", v); printf("\ngrf"); for (double v : grf.data) printf(" %.17g", v); printf("\n");
    printf("reset_msgs %d\n", mock::Bus::get().count["/gazebo/set_model_state"]);
    return 0;
}
