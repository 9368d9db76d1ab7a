#pragma once
#include <string>
#include <geometry_msgs/Pose.h>
namespace gazebo_msgs { struct ModelState { std::string model_name, reference_frame; geometry_msgs::Pose pose; geometry_msgs::Twist twist; }; }
