// ROS-free counterpart of the reference's src/whole_body_controller_node.cpp + WholeBodyController::controlLoop()
// (wbc.cpp:637-676) on the batched class: N robots in the nominal stance (BASELINE configs[0] when N = 1), tick 0 is the
// cold SQProblem::init, later ticks hot-start.  Prints per-tick solver statistics and the first robot's torques.
//   g++ -std=c++17 -Iinclude examples/whole_body_controller_node.cpp -Lquadrupedwholebodycontroller_b200 -lwbc_b200 -Wl,-rpath,... -o wbc_node
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "whole_body_controller_b200.hpp"

using anymal_wbc_b200::WholeBodyController;

int main(int argc, char** argv) {
    const int N = argc > 1 ? std::atoi(argv[1]) : 1, ticks = argc > 2 ? std::atoi(argv[2]) : 10;
    try {
        WholeBodyController wbc(N);                                   // == `WholeBodyController wbc;` of the reference node
        // what the three callbacks would deliver: nominal joints (wbc.cpp:81), base so that the feet touch z = 0, references = initialReferencePose
        const double q0[12] = {0.0, -0.4, 0.8, 0.0, 0.4, -0.8, 0.0, 0.4, -0.8, 0.0, -0.4, 0.8};
        std::vector<double> q(12 * N), qd(12 * N, 0.0), pos(3 * N, 0.0), quat(4 * N, 0.0), twist(6 * N, 0.0);
        std::vector<double> pose(6 * N, 0.0), vel(6 * N, 0.0), acc(6 * N, 0.0), sw(12 * N, 0.0);
        std::vector<unsigned char> contacts(4 * N, 1);
        for (int i = 0; i < N; i++) {
            for (int k = 0; k < 12; k++) q[k * N + i] = q0[k];
            pos[2 * N + i] = 0.585; quat[i] = 1.0; pose[2 * N + i] = 0.5;
        }
        std::vector<double> tau(12 * N), grf(12 * N); std::vector<int> rc(N), nwsr(N);
        int failures = 0;
        for (int t = 0; t < ticks; t++) {
            wbc.jointStateCallback(q.data(), qd.data());
            wbc.floatingBaseStateCallback(pos.data(), quat.data(), twist.data());
            wbc.referenceCallback(pose.data(), vel.data(), acc.data(), sw.data(), sw.data(), sw.data(), contacts.data());
            wbc.updateState(); wbc.solveQP(); wbc.computeJointTorques();     // wbc.cpp:650-652
            wbc.getOutputs(tau.data(), grf.data(), nullptr, rc.data(), nwsr.data(), nullptr, nullptr);
            double fz = 0; for (int l = 0; l < 4; l++) fz += grf[(3 * l + 2) * N];
            int bad = 0; long wsr = 0; for (int i = 0; i < N; i++) { bad += rc[i] != 0; wsr += nwsr[i]; }
            failures += bad;
            std::printf("tick %d: rc!=0 %d/%d  mean nWSR %.1f  robot0 sum f_z %.6f N  tau0 = [%.4f %.4f %.4f]\n", t, bad, N, double(wsr) / N, fz,
                        tau[0], tau[N], tau[2 * N]);
            if (bad && N == 1) { std::printf("QP did not solve!\n"); break; }   // wbc.cpp:654-659
        }
        return failures ? 2 : 0;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 1;
    }
}
