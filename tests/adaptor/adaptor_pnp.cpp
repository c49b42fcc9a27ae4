// adaptor_pnp.cpp -- the reference's simplePnP test (tests/test_sba_solver/main.cpp:61-236) written against
// include/sba_b200_adaptor.hpp: the blocks are plain double[3] owned by the caller and identified by address, the
// observations enter through ProjectorModule::addProjectionCostFunction, the solver follows init / opt_step /
// writeResults / cleanup.  Exit codes: 0 pass, 3 no CUDA device (init() returned false with the create error: the
// library has no CPU fallback), 1 any other failure.
//   part 1: exact observations, pose at identity -> cost < 1e-6 before and after (the reference's own assertion);
//   part 2: pose perturbed, points held constant, first run -> the LM brings the pose back (resection).
#include <cmath>
#include <cstdio>
#include <cstring>

#include "sba_b200_adaptor.hpp"

int main() {
    using namespace sba_b200;
    const double f = 35.0, pp[2] = {320.0, 240.0}, k[3] = {0, 0, 0}, t2[2] = {0, 0}, B[2] = {0, 0};
    double pts[5][3] = {{42, 69, 33}, {33, 27, 50}, {-6, -99, 70}, {-49, -28, 12}, {-69, -42, 33}};
    double uv[5][2];
    for (int i = 0; i < 5; i++) { uv[i][0] = f * pts[i][0] / pts[i][2] + pp[0]; uv[i][1] = f * pts[i][1] / pts[i][2] + pp[1]; }
    const double I2[4] = {1, 0, 0, 1};

    for (int part = 0; part < 2; part++) {
        double rAxis[3] = {0, 0, 0}, pos[3] = {0, 0, 0};
        if (part == 1) { rAxis[0] = 0.02; rAxis[1] = -0.015; rAxis[2] = 0.01; pos[0] = 0.3; pos[1] = -0.2; pos[2] = 0.25; }
        GpuSBASolver solver;
        solver.setOptimizationSteps(30);
        solver.setFuncTolerance(1e-14);
        solver.setParamsTolerance(1e-14);
        GpuPinholeProjectorModule cam = solver.addCameraModel(f, pp, k, t2, B);
        solver.registerPose(rAxis, pos, cam.model(), /*fixed*/ part == 0);
        double unknown[3] = {0, 0, 0};
        if (cam.addProjectionCostFunction(unknown, rAxis, pos, uv[0], I2)) {   // unregistered block: must be refused
            std::fprintf(stderr, "an unregistered point block was accepted\n");
            return 1;
        }
        {   // a foreign camera model, unregistered lever-arm blocks and a second offset on the same pose must be refused with
            // `false`, never forwarded as a plain observation
            double probe[3] = {1, 2, 30}, la_r[3] = {0, 0, 0}, la_t[3] = {0, 0, 0};
            solver.registerPoint(probe);
            if (cam.addProjectionCostFunction(probe, rAxis, pos, uv[0], I2, RigidOffset(), la_r, la_t)) { std::fprintf(stderr, "an unregistered lever arm was accepted\n"); return 1; }
            GpuPinholeProjectorModule other = solver.addCameraModel(2.0 * f, pp, k, t2, B);
            if (other.addProjectionCostFunction(probe, rAxis, pos, uv[0], I2)) { std::fprintf(stderr, "a pose of another camera model was accepted\n"); return 1; }
            if (solver.lastError().empty()) return 1;
        }
        for (int i = 0; i < 5; i++) {
            solver.registerPoint(pts[i], /*fixed*/ part == 1);
            if (!cam.addProjectionCostFunction(pts[i], rAxis, pos, uv[i], i % 2 ? I2 : nullptr)) {
                std::fprintf(stderr, "addProjectionCostFunction failed: %s\n", solver.lastError().c_str());
                return 1;
            }
        }
        if (!solver.init()) {
            std::fprintf(stderr, "init failed: %s\n", solver.lastError().c_str());
            return std::strstr(solver.lastError().c_str(), "device") || std::strstr(solver.lastError().c_str(), "CUDA") ? 3 : 1;
        }
        if (!solver.opt_step() || !solver.opt_step() /* later calls return true */ || !solver.writeResults()) {
            std::fprintf(stderr, "step failed: %s\n", solver.lastError().c_str());
            return 1;
        }
        const sba_summary& s = solver.summary();
        std::printf("part %d: initial cost %.3e final cost %.3e iterations %d pose r = (%.2e %.2e %.2e) t = (%.2e %.2e %.2e)\n", part,
                    s.initial_cost, s.final_cost, s.iterations, rAxis[0], rAxis[1], rAxis[2], pos[0], pos[1], pos[2]);
        if (part == 0 && !(s.initial_cost < 1e-6 && s.final_cost < 1e-6)) return 1;
        if (part == 1) {
            if (!(s.initial_cost > 1.0 && s.final_cost < 1e-12)) return 1;
            for (int a = 0; a < 3; a++) if (std::fabs(rAxis[a]) > 1e-7 || std::fabs(pos[a]) > 1e-6) return 1;
        }
        if (part == 1) {   // after the first observation of the pose (identity offset) another offset is refused
            RigidOffset off;
            off.t[0] = 0.1;
            if (cam.addProjectionCostFunction(pts[0], rAxis, pos, uv[0], I2, off)) { std::fprintf(stderr, "a second offset on one pose was accepted\n"); return 1; }
        }
        solver.cleanup();
    }
    {   // part 3: the PoseTransformLeverArmCost branch (pinholecameraprojectormodule.cpp:147-166) -- the same five points seen by a
        // sensor mounted on a body with a constant offset and a lever arm; body pose perturbed, points and lever arm constant:
        // the LM recovers the body pose that puts the SENSOR at the identity
        double rAxis[3] = {0.0, 0.0, 0.0}, pos[3] = {0.0, 0.0, 0.0}, la_r[3] = {0.0, 0.0, 0.0}, la_t[3] = {0.2, -0.1, 0.05};
        RigidOffset off;
        off.t[0] = 0.3; off.t[1] = 0.1; off.t[2] = -0.2;
        // sensor2world = (pose * offset) * lever = identity  <=>  body t = -(t_off + l) with all rotations zero
        const double truth[3] = {-(off.t[0] + la_t[0]), -(off.t[1] + la_t[1]), -(off.t[2] + la_t[2])};
        for (int a = 0; a < 3; a++) pos[a] = truth[a] + (a == 0 ? 0.2 : -0.1);
        rAxis[1] = 0.01;
        GpuSBASolver solver;
        solver.setOptimizationSteps(30);
        solver.setFuncTolerance(1e-14);
        solver.setParamsTolerance(1e-14);
        GpuPinholeProjectorModule cam = solver.addCameraModel(f, pp, k, t2, B);
        solver.registerPose(rAxis, pos, cam.model(), false);
        solver.registerLeverArm(la_r, la_t, /*fixed*/ true);
        for (int i = 0; i < 5; i++) {
            solver.registerPoint(pts[i], true);
            if (!cam.addProjectionCostFunction(pts[i], rAxis, pos, uv[i], I2, off, la_r, la_t)) {
                std::fprintf(stderr, "addProjectionCostFunction (offset + lever arm) failed: %s\n", solver.lastError().c_str());
                return 1;
            }
        }
        if (!solver.init()) {
            std::fprintf(stderr, "init failed: %s\n", solver.lastError().c_str());
            return std::strstr(solver.lastError().c_str(), "device") || std::strstr(solver.lastError().c_str(), "CUDA") ? 3 : 1;
        }
        if (!solver.opt_step() || !solver.writeResults()) { std::fprintf(stderr, "step failed: %s\n", solver.lastError().c_str()); return 1; }
        const sba_summary& s = solver.summary();
        std::printf("part 3: initial cost %.3e final cost %.3e pose t = (%.6f %.6f %.6f)\n", s.initial_cost, s.final_cost, pos[0], pos[1], pos[2]);
        if (!(s.initial_cost > 1.0 && s.final_cost < 1e-12)) return 1;
        for (int a = 0; a < 3; a++) if (std::fabs(rAxis[a]) > 1e-7 || std::fabs(pos[a] - truth[a]) > 1e-6) return 1;
    }
    std::puts("adaptor_pnp: ok");
    return 0;
}
