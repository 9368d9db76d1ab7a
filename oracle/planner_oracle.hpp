// ORACLE -- TEST INFRASTRUCTURE ONLY. Never linked into or called by the product path.
// Literal CPU restatement of the reference's MotionPlanner (src/motion_planner.cpp:5-97 Bezier / quintic helpers,
// :130-177 constructor, :180-376 plannerLoop; include/anymal_wbc/motion_planner.hpp; config/params_planner.yaml).
// The reference's plannerLoop is a blocking nested loop paced by ros::Rate(1/dt); here the SAME nested loop runs for a
// given number of planner periods: every `rate.sleep()` of the reference is one call of sleep() below, which ends the
// current period, records whether a message was published in it and reads the next /cmd_vel (ros::spinOnce).
// The reference package has no test, fixture or golden vector for the planner -> PARITY UNPINNED by the reference;
// the restatement keeps the reference's expression order (Eigen expressions evaluated coefficient-wise, left to right).
#pragma once
#include <vector>
#include <array>

namespace oracle {

// config/params_planner.yaml (body_initial_velocity is not in the yaml and never read by plannerLoop)
struct PlannerParams {
    double step_length=0.1, height_control_point=0.1, x_offset=0.50, y_offset=0.33, step_duration=0.2, body_height=0.50,
           body_final_velocity=0.40, dt=0.01;
};

// msg/WbcReferenceMsg.msg
struct ReferenceMsg {
    double desiredComPose[6], desiredComVelocity[6], desiredComAcceleration[6];
    double desiredSwingLegsPosition[12], desiredSwingLegsVelocity[12], desiredSwingLegsAcceleration[12];
    unsigned char footContacts[4];
};

struct PlannerRecord { bool published; ReferenceMsg msg; };     // state of ref_msg_ at the end of a planner period

// Runs the planner for cmd.size() periods; cmd[k] = (linear.x, linear.y, angular.z) in force DURING period k
// (i.e. delivered by the spinOnce that precedes it; period 0 sees cmd[0] as if it had arrived before the loop started).
std::vector<PlannerRecord> planner_run(const PlannerParams& p,const std::vector<std::array<double,3>>& cmd);

}  // namespace oracle
