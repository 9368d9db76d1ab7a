// ROS-free counterpart of running the reference's two nodes side by side -- src/motion_planner_node.cpp (MotionPlanner::run) and
// src/whole_body_controller_node.cpp (WholeBodyController::controlLoop, wbc.cpp:637-676) -- against a plant, for N robots on one GPU:
// every robot gets its own /cmd_vel, the planner runs at 100 Hz, the controller at 400 Hz, the ideal-actuator plant closes the loop.
//   g++ -std=c++17 -Iinclude examples/motion_planner_node.cpp -Lquadrupedwholebodycontroller_b200 -lwbc_b200 -Wl,-rpath,... -o planner_node
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "whole_body_controller_b200.hpp"

using anymal_wbc_b200::MotionPlanner;
using anymal_wbc_b200::WholeBodyController;

int main(int argc, char** argv) {
    const int N = argc > 1 ? std::atoi(argv[1]) : 4, ticks = argc > 2 ? std::atoi(argv[2]) : 400;
    try {
        WholeBodyController wbc(N);
        MotionPlanner planner(wbc);
        wbc.enablePlant();
        std::vector<double> vx(N), vy(N, 0.0), wz(N, 0.0);
        for (int i = 0; i < N; i++) { vx[i] = 0.1 + 0.3 * double(i % 8) / 7.0; if (i % 4 == 3) wz[i] = 0.2; }
        planner.input_callback(vx.data(), vy.data(), wz.data());
        std::vector<double> pos(3 * N), com(3 * N); std::vector<int> n(N), bad(N), wsr(N), wmax(N);
        const auto t0 = std::chrono::steady_clock::now();
        for (int done = 0; done < ticks; done += 100) {
            const int chunk = ticks - done < 100 ? ticks - done : 100;
            wbc.rollout(chunk);                                              // planner period every 4th tick, then the three hot methods, then the plant
            wbc.getState(nullptr, nullptr, pos.data(), nullptr, nullptr); wbc.getCenterOfMass(com.data());
            wbc.getRolloutStatistics(n.data(), bad.data(), wsr.data(), wmax.data());
            long f = 0, s = 0; for (int i = 0; i < N; i++) { f += bad[i]; s += wsr[i]; }
            std::printf("tick %d: robot0 base (%.4f %.4f %.4f) com z %.4f  failed solves so far %ld  mean nWSR %.2f\n", done + chunk, pos[0], pos[N], pos[2 * N],
                        com[2 * N], f, double(s) / (double(N) * (done + chunk)));
        }
        wbc.synchronize();
        const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        long f = 0; for (int i = 0; i < N; i++) f += bad[i];
        std::printf("%d robots x %d ticks in %.3f s -> %.0f solves/s, failed solves %ld\n", N, ticks, secs, double(N) * ticks / secs, f);
        return f ? 2 : 0;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 1;
    }
}
