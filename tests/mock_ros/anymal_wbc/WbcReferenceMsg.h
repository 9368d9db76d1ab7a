// msg/WbcReferenceMsg.msg:1-7 of the reference as the struct genmsg would emit
#pragma once
#include <array>
#include <std_msgs/Float64MultiArray.h>
namespace anymal_wbc { struct WbcReferenceMsg { std_msgs::Float64MultiArray desiredComPose, desiredComVelocity, desiredComAcceleration, desiredSwingLegsPosition, desiredSwingLegsVelocity, desiredSwingLegsAcceleration; std::array<unsigned char, 4> footContacts{{1, 1, 1, 1}}; }; }
