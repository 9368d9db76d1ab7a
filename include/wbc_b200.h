/* wbc_b200.h -- C ABI of the batched ANYmal whole-body controller for NVIDIA B200 (sm_100a).
 *
 * Drop-in boundary for ONE hot path of Mavagnale/QuadrupedWholeBodyController: the body of the 400 Hz control loop
 *     updateState(); solveQP(); computeJointTorques();          (reference src/whole_body_controller.cpp:650-652)
 * evaluated for a batch of N independent robots on one GPU.  Every entry point below names the reference interface it
 * replaces (paths relative to the reference repository).  There is no CPU fallback: every call fails with
 * WBC_ERR_CUDA if no sm_100-class device is usable.
 *
 * Array convention: structure-of-arrays, instance index fastest -- element k of robot i is a[k*N + i]; fp64; flags are
 * bytes.  Pointers may be HOST or DEVICE pointers (detected with cudaPointerGetAttributes); host buffers are copied
 * with cudaMemcpyAsync on the handle's stream.  BUFFER LIFETIME: input copies are asynchronous -- from PINNED host memory and
 * from device memory the caller's buffer is read when the stream reaches the copy, so it must stay valid and unmodified until
 * then (wbc_synchronize, or any later call that returns results to a host pointer); pageable host memory is staged by the
 * driver before the call returns.  Output copies to host pointers are complete when the getter returns.  The caller owns its buffers, the library owns all per-robot
 * persistent state (finite-difference history, integral error, QP warm-start state) in device memory.
 * Threading: one handle = one host thread at a time; handles are independent.
 *
 * Orders fixed by the reference: joints/DOFs LH_HAA,LH_HFE,LH_KFE, LF_*, RF_*, RH_* (whole_body_controller.cpp:81,234;
 * config/controllers.yaml:8-19); legs / contacts / ground-reaction forces LH, LF, RF, RH (:327-341);
 * QP variables x[42] = [CoM acc(6) | joint acc(12) | contact forces(12) | slacks(12)] (whole_body_controller.hpp:31).
 */
#ifndef WBC_B200_H
#define WBC_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define WBC_NUM_JOINTS 12          /* whole_body_controller.hpp:27 numberOfJoints          */
#define WBC_NUM_LEGS 4             /* whole_body_controller.hpp:28 numberOfLegs            */
#define WBC_QP_NV 42               /* whole_body_controller.hpp:31 qpNumberOfVariables     */
#define WBC_QP_NC 70               /* whole_body_controller.hpp:32 qpNumberOfConstraints   */

#define WBC_OK 0
#define WBC_ERR_ARG (-1)
#define WBC_ERR_CUDA (-2)
#define WBC_ERR_NOMEM (-3)

/* per-robot solver status values are qpOASES returnValue numbers (lib/qpOASES/include/qpOASES/MessageHandling.hpp):
 * 0 SUCCESSFUL_RETURN, 64 RET_MAX_NWSR_REACHED, 37 RET_INIT_FAILED_INFEASIBILITY, 61 RET_HOTSTART_STOPPED_INFEASIBILITY ... */

typedef struct wbc_handle wbc_handle;

/* Replaces WholeBodyController::params / loadParameters (whole_body_controller.hpp:91-104, whole_body_controller.cpp:122-148,
 * config/params_controller.yaml).  Shared by the whole batch. */
typedef struct wbc_params {
    double friction, loopRate, maxTorque, kpValue, kpValueZ, kdValue, kiValue, kpSwingValue, kdSwingValue, slackWeight;
    double initialReferencePose[6];
} wbc_params;

/* config/params_controller.yaml:1-11 */
void wbc_default_params(wbc_params* p);

/* Replaces WholeBodyController::WholeBodyController() (whole_body_controller.cpp:22-59): model "load" (the 13-body ANYmal
 * table compiled from urdf/anymal.urdf is baked into the library), SQProblem(42,70) construction, loadParameters,
 * setInitialState.  device = CUDA ordinal.  stream = a cudaStream_t (as void*) to run on, or NULL for a private stream. */
int wbc_create(int n_robots, int device, const wbc_params* params, void* stream, wbc_handle** out);
/* (params->slackWeight < 1 is refused with WBC_ERR_ARG: the solver leaves out qpOASES' flipping-bound and zero-curvature branches,
 * QProblem.cpp:3938-3963, 4278-4511, which cannot fire while H >= I -- see the header of csrc/wbc_qp.cuh.) */
int wbc_destroy(wbc_handle* h);                                   /* ~WholeBodyController (whole_body_controller.cpp:61-63) */

/* Replaces setInitialState() (whole_body_controller.cpp:65-120) for robots [first, first+count): contacts 1111, nominal joint
 * configuration, base at z = 0.60, references = initialReferencePose, old T = I, old Jacobians = 0, integral error = 0,
 * firstControllerIteration_ = true (next solve is SQProblem::init).  Tdot^-1 (uninitialised upstream) is defined as 0. */
int wbc_reset(wbc_handle* h, int first, int count);

/* Replaces jointStateCallback (whole_body_controller.cpp:232-254; q, qd already in model DOF order) and
 * floatingBaseStateCallback (:187-230): base position[3], orientation quaternion (w,x,y,z)[4] (Eigen toRotationMatrix, not
 * normalised), base twist[6] = [linear world; angular world]. */
int wbc_set_state(wbc_handle* h, const double* q, const double* qd, const double* base_pos, const double* base_quat_wxyz,
                  const double* base_twist);

/* Replaces referenceCallback (whole_body_controller.cpp:150-185; msg/WbcReferenceMsg.msg): desired CoM pose[6] = [xyz; rpy],
 * velocity[6], acceleration[6], swing-leg position/velocity/acceleration[12], footContacts[4] (bytes, leg-major: c[l*N+i]).
 * switching[N] (bytes) = isSwitchingFootState_; NULL derives it from the previous contacts exactly like :176-184. */
int wbc_set_reference(wbc_handle* h, const double* pose, const double* vel, const double* acc, const double* sw_pos,
                      const double* sw_vel, const double* sw_acc, const unsigned char* contacts, const unsigned char* switching);

/* The three hot methods, batched (whole_body_controller.hpp:47,54,52; whole_body_controller.cpp:256-294, 466-542, 553-577).
 * A robot whose solve did not return 0 gets zero torques for that tick (what terminate() publishes, :626-635); the next call
 * hot-starts it again exactly like the SQProblem object would -- halting the loop (:654-659) is the caller's decision. */
int wbc_update_state(wbc_handle* h);
int wbc_solve_qp(wbc_handle* h);
int wbc_compute_joint_torques(wbc_handle* h);   /* the recovery itself runs as the epilogue of the solve kernel (x is still on chip); this call orders after it */
/* No counterpart in the reference.  solveQP() for callers that must not allocate (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc): wbc_solve_qp allocates its 22 KB-per-robot warm-start hand-over buffer
 * with cudaMalloc at the first hot-started tick; this variant launches the same kernels with that scratch (and the work counter) in a file-scope
 * static __device__ array of the library, and calls no cudaMalloc / cudaFree / allocating routine.  Limit: handles of at most
 * WBC_STATIC_MAX_ROBOTS = 1024 robots (22.5 MB held by every GPU that loads the library until exit); a larger handle is refused with
 * WBC_ERR_ARG and a text in wbc_last_error.  Inputs and outputs are the caller's, exactly as for wbc_solve_qp (wbc_set_state / wbc_set_reference
 * + wbc_update_state before, wbc_get_outputs after -- none of them allocates); results are bit-identical to wbc_solve_qp.  ALL calls on one
 * device share the array, which keeps the last call's contents (the call clears the one word that must start at zero): calls that overlap on
 * the GPU -- two handles on different streams of one device -- race; serialise them. */
#define WBC_STATIC_MAX_ROBOTS 1024
int wbc_solve_qp_static(wbc_handle* h);
int wbc_step(wbc_handle* h);                                     /* the three calls above, asynchronous on the stream */
int wbc_synchronize(wbc_handle* h);

/* Replaces the publishers (whole_body_controller.cpp:558-576, 666-670) and quadraticProblem_ getters: joint torques tau[12],
 * desired ground reaction forces grf[12], x[42], qpReturnValue_ rc[N], working-set recalculations nwsr[N], working set
 * ws[112] in {-1,0,+1} (QProblem::getWorkingSet: 42 bounds then 70 constraints), CoM position com[3].  Any pointer may be NULL.
 * Synchronises the stream when a HOST pointer is given. */
int wbc_get_outputs(wbc_handle* h, double* tau, double* grf, double* x, int* rc, int* nwsr, signed char* ws, double* com);

/* QProblem::getDualSolution of every robot's SQProblem (lib/qpOASES/src/QProblemB.cpp:800-823): y[112] = 42 bound multipliers then 70
 * constraint multipliers, qpOASES sign convention (>= 0 at an active lower side, <= 0 at an active upper side).  Host or device. */
int wbc_get_duals(wbc_handle* h, double* y);
/* 1 while hot-started ticks run the TQ rebuild as a kernel of its own (needs 22 KB of device memory per robot); 0 after that
 * allocation failed and the rebuild moved inside the solve kernel (identical results, slower) -- wbc_last_error says so once. */
int wbc_warm_start_split(const wbc_handle* h);
/* Explicit switches for tests and measurements; every setting gives bit-identical results.  Keys: "warm_start_split" (0 = rebuild the
 * TQ factorisation inside the solve kernel), "cohort" (-1 auto, 0 never, 1 always: gang scheduling of the solve kernel),
 * "setup_canary" (1 = guard the workspace prefix of the warm-start kernel), "rollout_graph" (0 = wbc_rollout never replays a CUDA graph),
 * "park" (0 = an all-cold batch runs as one launch of the general solve kernel instead of the two launches that park and resume solves; the parking buffer is 40 KB per
 * robot of a chunk of at most 65536 robots, allocated at the first all-cold solve), "park_chunk" (robots per chunk; before the first cold solve). */
int wbc_set_option(wbc_handle* h, const char* key, int value);

/* Test / inspection hook: the dense QP of robot i as the reference would hand it to qpOASES (row-major H[42*42], g[42],
 * A[70*42], lbA[70], ubA[70]; whole_body_controller.cpp:479-515), expanded on the host from the compact device form. */
int wbc_get_dense_qp(wbc_handle* h, int robot, double* H, double* g, double* A, double* lbA, double* ubA);

/* ------------------------------------------------------------------------------------------------------------------------------
 * "Next" rows of the hot path: the producer of the references and a plant, so that a closed-loop
 * rollout of N robots never leaves the GPU.
 * ------------------------------------------------------------------------------------------------------------------------------ */

/* Replaces MotionPlanner::params / load_parameters (include/anymal_wbc/motion_planner.hpp:40-51, src/motion_planner.cpp:99-119,
 * config/params_planner.yaml).  cycle_duration = 4 * step_duration is derived like :118. */
typedef struct wbc_planner_params {
    double step_length, height_control_point, x_offset, y_offset, step_duration, body_height, body_final_velocity, dt;
} wbc_planner_params;
void wbc_planner_default_params(wbc_planner_params* p);          /* config/params_planner.yaml:1-8 */

/* Replaces MotionPlanner::MotionPlanner() (src/motion_planner.cpp:130-177): one planner per robot of the handle, velocity command
 * zero, body at (0,0,body_height), feet at (+-x_offset, +-y_offset, 0), latched message = stand at body_height, contacts 1111. */
int wbc_planner_create(wbc_handle* h, const wbc_planner_params* params);
int wbc_planner_reset(wbc_handle* h, int first, int count);
/* Replaces MotionPlanner::input_callback (src/motion_planner.cpp:122-127; geometry_msgs/Twist linear.x, linear.y, angular.z):
 * arrays of N, host or device; NULL keeps the previous values.  Takes effect at the next planner period (ros::spinOnce). */
int wbc_planner_set_cmd_vel(wbc_handle* h, const double* vx, const double* vy, const double* yaw_rate);
/* One period (params.dt) of MotionPlanner::plannerLoop (src/motion_planner.cpp:180-376) for every robot.  Where the reference
 * publishes /anymal/reference in that period, the message is handed to the controller exactly like referenceCallback
 * (src/whole_body_controller.cpp:150-185: all six reference vectors, footContacts, isSwitchingFootState_ re-derived from the previous
 * message); where it does not (phase change, end of cycle), the controller's reference and switching flag stay latched. */
int wbc_planner_tick(wbc_handle* h);
/* Inspection: the controller's current reference buffers (what referenceCallback last stored), the switching flag and whether the
 * last planner period published (bytes; needs wbc_planner_create).  Any pointer may be NULL. */
int wbc_get_reference(wbc_handle* h, double* pose, double* vel, double* acc, double* sw_pos, double* sw_vel, double* sw_acc,
                      unsigned char* contacts, unsigned char* switching, unsigned char* published);
/* Inspection: the controller's current state inputs (what the state callbacks last stored, or what the plant integrated). */
int wbc_get_state(wbc_handle* h, double* q, double* qd, double* base_pos, double* base_quat_wxyz, double* base_twist);

/* Plant for closed-loop rollouts.  The reference has none (Gazebo, launch/spawn.launch:27-59); this is an ideal-actuator rigid-body
 * plant: it integrates the generalized accelerations implied by the QP solution (its equality rows ARE the floating-base dynamics
 * and the stance constraints), nu_dot = T^-1 (x[0:18] - Tdot nu), with semi-implicit Euler at loopRate and writes the new
 * q, qd, base pose and twist into the controller's state inputs.  wbc_plant_enable makes wbc_update_state export T^-1 / Tdot nu. */
int wbc_plant_enable(wbc_handle* h);
int wbc_plant_step(wbc_handle* h);
/* `ticks` control ticks of the closed loop (src/whole_body_controller.cpp:639-673 with the planner node beside it):
 * every `planner_every`-th tick (4 = 400 Hz / 100 Hz) one planner period, then updateState, solveQP, computeJointTorques, plant step.
 * Asynchronous on the stream.  Needs wbc_planner_create and wbc_plant_enable.  In the steady state (every robot hot-started) a whole planner
 * period is replayed as ONE CUDA graph launch (captured on first use; wbc_set_option "rollout_graph" 0 switches it off): same kernels in
 * the same order, same results; the per-kernel timers of wbc_last_kernel_ms are only updated by ticks issued as plain launches. */
int wbc_rollout(wbc_handle* h, int ticks, int planner_every);
/* per-robot rollout statistics since the last reset: control ticks, ticks whose solve did not return 0, sum and max of nWSR */
int wbc_rollout_stats(wbc_handle* h, int* ticks, int* failed_ticks, int* nwsr_sum, int* nwsr_max, int reset);

/* Verification build only (libwbc_b200_exact.so, -DWBC_EXACT -fmad=false: qpOASES' arithmetic operation by operation, bit-identical iterates):
 * out3 = { full linear-independence tests run (QProblem.cpp:2998-3072, 3500-3570), tests inside the undecided band of the product build's shortcut,
 * tests whose verdict the shortcut would have got wrong }.  Other builds return WBC_ERR_ARG. */
int wbc_debug_counters(wbc_handle* h, unsigned long long* out3, int reset);
int wbc_num_robots(const wbc_handle* h);
/* number of kernel launches issued by this handle so far (bench.py reports it as gpu_launches) */
long long wbc_launch_count(const wbc_handle* h);
/* duration [ms] of the most recent wbc_solve_qp kernel measured with CUDA events on the handle's stream (synchronises) */
int wbc_last_kernel_ms(wbc_handle* h, float* rbd_ms, float* qp_ms, float* torque_ms);
/* measurement hook: per-robot record of the last solve, 10 int64 per robot: clock64 ticks in data shift, step direction, step,
 * active-set change, projected Cholesky, ramp/drift, whole solve; iterations; ALGORITHMIC flops of the qpOASES method on dense
 * data (formulas of SURVEY.md section 8d, the numerator of the FP64 roofline); cold-start flag.  enable != 0 allocates the buffer.
 * Only in the profiling build of the same library (libwbc_b200_prof.so, -DWBC_PROFILE); the product build returns WBC_ERR_ARG. */
int wbc_profile(wbc_handle* h, int enable, long long* host_out);
/* measured FP64 FMA peak of the device in GFLOP/s (8 independent DFMA chains per thread, all SMs) -- the FP64 roofline denominator */
int wbc_measure_fp64_peak(wbc_handle* h, double* gflops);
/* measured shared-memory load bandwidth of the device in GB/s (conflict-free 64-bit loads, all SMs) -- the denominator for the solver's
 * shared-memory traffic (SURVEY.md section 8d asks for a measured figure, not a datasheet one) */
int wbc_measure_smem_peak(wbc_handle* h, double* gbytes_per_s);
const char* wbc_last_error(const wbc_handle* h);

/* ------------------------------------------------------------------------------------------------------------------------------
 * One batch over several GPUs of the box.  The batch shards by robot index with no exchange on the hot path:
 * device k owns the contiguous range [first_k, first_k + count_k) (remainder to the lowest shards) for the lifetime of the object, with its
 * own stream; every call below is issued from the calling host thread and is asynchronous on those streams unless it says otherwise.
 * Arrays are structure-of-arrays over the WHOLE batch (a[j*n_robots + i]), in host memory or in device memory of any GPU of the box.
 * ------------------------------------------------------------------------------------------------------------------------------ */
typedef struct wbc_multi wbc_multi;
/* Replaces N x WholeBodyController::WholeBodyController() (whole_body_controller.cpp:22-59) spread over n_devices GPUs. */
int wbc_multi_create(int n_robots, const int* device_ids, int n_devices, const wbc_params* params, wbc_multi** out);
int wbc_multi_destroy(wbc_multi* m);
int wbc_multi_num_devices(const wbc_multi* m);
int wbc_multi_num_robots(const wbc_multi* m);
int wbc_multi_shard(const wbc_multi* m, int k, int* device, int* first, int* count);
wbc_handle* wbc_multi_handle(wbc_multi* m, int k);               /* the shard's own handle, for everything not wrapped here (planner, plant, rollout) */
int wbc_multi_reset(wbc_multi* m);                                /* setInitialState (whole_body_controller.cpp:65-120) for every robot */
/* The three callbacks (whole_body_controller.cpp:150-254) for the whole batch: one strided copy per array and device. */
int wbc_multi_set_state(wbc_multi* m, const double* q, const double* qd, const double* base_pos, const double* base_quat_wxyz, const double* base_twist);
int wbc_multi_set_reference(wbc_multi* m, const double* pose, const double* vel, const double* acc, const double* sw_pos, const double* sw_vel,
                            const double* sw_acc, const unsigned char* contacts, const unsigned char* switching);
int wbc_multi_step(wbc_multi* m);                                 /* updateState; solveQP; computeJointTorques (whole_body_controller.cpp:650-652) on every device */
int wbc_multi_synchronize(wbc_multi* m);
/* Publishers (whole_body_controller.cpp:558-576, 666-670) of every shard into global arrays.  With DEVICE pointers of one GPU this is the optional
 * single gather of the results over NVLink (about 200 B per robot; peer-to-peer where the topology allows).  Blocks until the copies are done. */
int wbc_multi_get_outputs(wbc_multi* m, double* tau, double* grf, double* x, int* rc, int* nwsr, signed char* ws, double* com);
int wbc_multi_last_gather_ms(wbc_multi* m, float* ms);            /* device time of those copies, max over the GPUs */
const char* wbc_multi_last_error(const wbc_multi* m);

#ifdef __cplusplus
}
#endif
#endif /* WBC_B200_H */
