// whole_body_controller_b200.hpp -- ROS-free C++ mirror of the reference's WholeBodyController for a BATCH of robots.
//
// Same class name, constants and hot-method names as include/anymal_wbc/whole_body_controller.hpp:27-63 of the
// reference, so code written against the reference class reads the same; everything numerical is delegated to the CUDA
// library through the C ABI (wbc_b200.h).  Header-only: link with -lwbc_b200.  No CPU fallback: the constructor throws
// std::runtime_error if the library cannot use the requested CUDA device.
//
//   reference member / method                               here
//   ------------------------------------------------------  ---------------------------------------------------------
//   WholeBodyController()            (wbc.cpp:22-59)        WholeBodyController(nRobots, device, Params, stream)
//   loadParameters()                 (wbc.cpp:122-148)      Params (same field names as the `params` struct, wbc.hpp:91-104)
//   setInitialState()                (wbc.cpp:65-120)       setInitialState(first, count)
//   jointStateCallback(msg)          (wbc.cpp:232-254)      jointStateCallback(q, qd)             [DOF order LH,LF,RF,RH x HAA,HFE,KFE]
//   floatingBaseStateCallback(msg)   (wbc.cpp:187-230)      floatingBaseStateCallback(pos, quat_wxyz, twist)
//   referenceCallback(msg)           (wbc.cpp:150-185)      referenceCallback(pose, vel, acc, swPos, swVel, swAcc, contacts[, switching])
//   updateState/solveQP/computeJointTorques (wbc.cpp:256-577)   same names, whole batch, asynchronous on the stream
//   jointTorquePub_ / desiredGroundReactionForcesPub_ / centerOfMassPub_ (wbc.cpp:558-576,666-670)   getJointTorques / getGroundReactionForces / getCenterOfMass
//   qpSolution_, qpReturnValue_      (wbc.hpp:165-166)      getQpSolution, getQpReturnValues, getWorkingSetRecalculations, getWorkingSet
//   controlLoop() body               (wbc.cpp:650-652)      step()
//
// Array convention of every pointer argument: structure-of-arrays, robot index fastest (a[k*N + i]), host or device memory.
#ifndef WHOLE_BODY_CONTROLLER_B200_HPP
#define WHOLE_BODY_CONTROLLER_B200_HPP

#include <stdexcept>
#include <string>
#include <vector>

#include "wbc_b200.h"

namespace anymal_wbc_b200 {

const int numberOfJoints = WBC_NUM_JOINTS;                 // wbc.hpp:28
const int numberOfLegs = WBC_NUM_LEGS;                     // wbc.hpp:29
const double gravityAcceleration = 9.81;                   // wbc.hpp:30
const int qpNumberOfVariables = 6 + numberOfJoints + 3 * numberOfLegs + 3 * numberOfLegs;                                   // wbc.hpp:31
const int qpNumberOfConstraints = 6 + 3 * numberOfLegs + 4 * numberOfLegs + numberOfJoints + 6 * numberOfLegs;              // wbc.hpp:32
static_assert(qpNumberOfVariables == WBC_QP_NV && qpNumberOfConstraints == WBC_QP_NC, "QP dimensions of the reference");

class WholeBodyController {
public:
    struct Params : wbc_params {                          // wbc.hpp:91-104; defaults = config/params_controller.yaml
        Params() { wbc_default_params(this); }
    };

    explicit WholeBodyController(int nRobots = 1, int device = 0, const Params& params = Params(), void* cudaStream = nullptr)
        : n_(nRobots) {
        const int rc = wbc_create(nRobots, device, &params, cudaStream, &h_);
        if (rc != WBC_OK) throw std::runtime_error("WholeBodyController: wbc_create failed with code " + std::to_string(rc) +
                                                   " (no usable sm_100a CUDA device? there is no CPU fallback)");
    }
    ~WholeBodyController() { if (h_) wbc_destroy(h_); }
    WholeBodyController(const WholeBodyController&) = delete;
    WholeBodyController& operator=(const WholeBodyController&) = delete;

    int numberOfRobots() const { return n_; }

    void setInitialState() { check(wbc_reset(h_, 0, n_)); }
    void setInitialState(int first, int count) { check(wbc_reset(h_, first, count)); }

    void jointStateCallback(const double* jointPos, const double* jointVel) { check(wbc_set_state(h_, jointPos, jointVel, nullptr, nullptr, nullptr)); }
    void floatingBaseStateCallback(const double* basePos, const double* baseQuatWXYZ, const double* baseTwist) {
        check(wbc_set_state(h_, nullptr, nullptr, basePos, baseQuatWXYZ, baseTwist));
    }
    void referenceCallback(const double* desiredPose, const double* desiredCoMVelocity, const double* desiredCoMAcceleration,
                           const double* desiredSwingLegsPosition, const double* desiredSwingLegsVelocity, const double* desiredSwingLegsAcceleration,
                           const unsigned char* footContacts, const unsigned char* isSwitchingFootState = nullptr) {
        check(wbc_set_reference(h_, desiredPose, desiredCoMVelocity, desiredCoMAcceleration, desiredSwingLegsPosition, desiredSwingLegsVelocity,
                                desiredSwingLegsAcceleration, footContacts, isSwitchingFootState));
    }

    void updateState() { check(wbc_update_state(h_)); }
    void solveQP() { check(wbc_solve_qp(h_)); }
    void computeJointTorques() { check(wbc_compute_joint_torques(h_)); }
    void step() { check(wbc_step(h_)); }                    // the body of controlLoop(), wbc.cpp:650-652
    void synchronize() { check(wbc_synchronize(h_)); }

    void getJointTorques(double* tau) { check(wbc_get_outputs(h_, tau, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr)); }
    void getGroundReactionForces(double* grf) { check(wbc_get_outputs(h_, nullptr, grf, nullptr, nullptr, nullptr, nullptr, nullptr)); }
    void getQpSolution(double* x) { check(wbc_get_outputs(h_, nullptr, nullptr, x, nullptr, nullptr, nullptr, nullptr)); }
    void getQpReturnValues(int* rc) { check(wbc_get_outputs(h_, nullptr, nullptr, nullptr, rc, nullptr, nullptr, nullptr)); }
    void getWorkingSetRecalculations(int* nWSR) { check(wbc_get_outputs(h_, nullptr, nullptr, nullptr, nullptr, nWSR, nullptr, nullptr)); }
    void getWorkingSet(signed char* ws) { check(wbc_get_outputs(h_, nullptr, nullptr, nullptr, nullptr, nullptr, ws, nullptr)); }
    void getCenterOfMass(double* com) { check(wbc_get_outputs(h_, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, com)); }
    void getOutputs(double* tau, double* grf, double* x, int* rc, int* nWSR, signed char* ws, double* com) {
        check(wbc_get_outputs(h_, tau, grf, x, rc, nWSR, ws, com));
    }
    // the dense QP of one robot exactly as the reference hands it to qpOASES (wbc.cpp:479-515), row-major
    void getDenseQP(int robot, double* H, double* g, double* A, double* lbA, double* ubA) { check(wbc_get_dense_qp(h_, robot, H, g, A, lbA, ubA)); }

    // state inputs as last delivered by the callbacks or integrated by the plant; reference as last stored by referenceCallback
    void getState(double* q, double* qd, double* basePos, double* baseQuatWXYZ, double* baseTwist) { check(wbc_get_state(h_, q, qd, basePos, baseQuatWXYZ, baseTwist)); }
    void getReference(double* pose, double* vel, double* acc, double* swPos, double* swVel, double* swAcc, unsigned char* contacts, unsigned char* switching) {
        check(wbc_get_reference(h_, pose, vel, acc, swPos, swVel, swAcc, contacts, switching, nullptr));
    }
    // closed loop on the device (needs a MotionPlanner attached to this controller): plant + `ticks` control ticks, one planner period
    // every `plannerEvery` ticks (400 Hz loop, wbc.cpp:639-673; 100 Hz planner, motion_planner.cpp:182)
    void enablePlant() { check(wbc_plant_enable(h_)); }
    void plantStep() { check(wbc_plant_step(h_)); }
    void rollout(int ticks, int plannerEvery = 4) { check(wbc_rollout(h_, ticks, plannerEvery)); }
    void getRolloutStatistics(int* ticks, int* failedTicks, int* nwsrSum, int* nwsrMax, bool reset = false) {
        check(wbc_rollout_stats(h_, ticks, failedTicks, nwsrSum, nwsrMax, reset ? 1 : 0));
    }

    wbc_handle* handle() { return h_; }

private:
    void check(int rc) const {
        if (rc != WBC_OK) throw std::runtime_error(std::string("WholeBodyController: ") + wbc_last_error(h_) + " (code " + std::to_string(rc) + ")");
    }
    wbc_handle* h_ = nullptr;
    int n_ = 0;
};

// The same batch spread over several GPUs of the box from ONE host thread (wbc_multi_*): device k owns a contiguous range of robots with its own
// stream, no exchange on the hot path.  Pointers are arrays over the WHOLE batch, host memory or device memory of any GPU; getOutputs
// into device memory of one GPU is the optional single gather of the results over NVLink (lastGatherMilliseconds() = its device time).
class MultiGpuWholeBodyController {
public:
    MultiGpuWholeBodyController(int nRobots, const std::vector<int>& devices, const WholeBodyController::Params& params = WholeBodyController::Params()) : n_(nRobots) {
        const int rc = wbc_multi_create(nRobots, devices.data(), (int)devices.size(), &params, &m_);
        if (rc != WBC_OK) throw std::runtime_error("MultiGpuWholeBodyController: wbc_multi_create failed with code " + std::to_string(rc));
    }
    ~MultiGpuWholeBodyController() { if (m_) wbc_multi_destroy(m_); }
    MultiGpuWholeBodyController(const MultiGpuWholeBodyController&) = delete;
    MultiGpuWholeBodyController& operator=(const MultiGpuWholeBodyController&) = delete;
    int numberOfRobots() const { return n_; }
    int numberOfDevices() const { return wbc_multi_num_devices(m_); }
    void shard(int k, int& device, int& first, int& count) const { check(wbc_multi_shard(m_, k, &device, &first, &count)); }
    void setInitialState() { check(wbc_multi_reset(m_)); }
    void jointStateCallback(const double* jointPos, const double* jointVel) { check(wbc_multi_set_state(m_, jointPos, jointVel, nullptr, nullptr, nullptr)); }
    void floatingBaseStateCallback(const double* basePos, const double* baseQuatWXYZ, const double* baseTwist) { check(wbc_multi_set_state(m_, nullptr, nullptr, basePos, baseQuatWXYZ, baseTwist)); }
    void referenceCallback(const double* pose, const double* vel, const double* acc, const double* swPos, const double* swVel, const double* swAcc,
                           const unsigned char* footContacts, const unsigned char* isSwitchingFootState = nullptr) {
        check(wbc_multi_set_reference(m_, pose, vel, acc, swPos, swVel, swAcc, footContacts, isSwitchingFootState));
    }
    void step() { check(wbc_multi_step(m_)); }               // updateState(); solveQP(); computeJointTorques(); on every device
    void synchronize() { check(wbc_multi_synchronize(m_)); }
    void getOutputs(double* tau, double* grf, double* x, int* rc, int* nWSR, signed char* ws, double* com) { check(wbc_multi_get_outputs(m_, tau, grf, x, rc, nWSR, ws, com)); }
    float lastGatherMilliseconds() { float t = 0; check(wbc_multi_last_gather_ms(m_, &t)); return t; }
    wbc_handle* shardHandle(int k) { return wbc_multi_handle(m_, k); }

private:
    void check(int rc) const { if (rc != WBC_OK) throw std::runtime_error(std::string("MultiGpuWholeBodyController: ") + wbc_multi_last_error(m_) + " (code " + std::to_string(rc) + ")"); }
    wbc_multi* m_ = nullptr;
    int n_ = 0;
};

// Batched mirror of the reference's MotionPlanner (include/anymal_wbc/motion_planner.hpp:10-52, src/motion_planner.cpp): one planner
// per robot of a WholeBodyController, living on the same GPU.  plannerLoop() is a blocking loop paced by ros::Rate in the
// reference; here one call of plannerPeriod() is one period (params.dt) of that loop for the whole batch, and a published
// message reaches the controller exactly like referenceCallback.
class MotionPlanner {
public:
    struct Params : wbc_planner_params {                   // motion_planner.hpp:40-51; defaults = config/params_planner.yaml
        Params() { wbc_planner_default_params(this); }
    };
    explicit MotionPlanner(WholeBodyController& controller, const Params& params = Params()) : h_(controller.handle()) {
        if (wbc_planner_create(h_, &params) != WBC_OK) throw std::runtime_error(std::string("MotionPlanner: ") + wbc_last_error(h_));
    }
    // input_callback (motion_planner.cpp:122-127): linear.x, linear.y, angular.z of /cmd_vel, arrays of nRobots (host or device), nullptr keeps
    void input_callback(const double* linearX, const double* linearY, const double* angularZ) { check(wbc_planner_set_cmd_vel(h_, linearX, linearY, angularZ)); }
    void plannerPeriod() { check(wbc_planner_tick(h_)); }   // one pass of the pacing of plannerLoop (motion_planner.cpp:180-376)
    void reset() { check(wbc_planner_reset(h_, 0, wbc_num_robots(h_))); }
    void getPublished(unsigned char* published) { check(wbc_get_reference(h_, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, published)); }

private:
    void check(int rc) const { if (rc != WBC_OK) throw std::runtime_error(std::string("MotionPlanner: ") + wbc_last_error(h_) + " (code " + std::to_string(rc) + ")"); }
    wbc_handle* h_;
};

}  // namespace anymal_wbc_b200
#endif
