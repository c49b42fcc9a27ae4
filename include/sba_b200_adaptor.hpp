// sba_b200_adaptor.hpp -- header-only C++ adaptor that gives the C ABI of
// sba_b200.h the shape of the reference's own interfaces, so that a steviapp
// build can route ModularSBASolver's camera/point path to the GPU engine
// without touching its SBA modules.
//
// Mirrors (names, argument meaning, bool error convention):
//   StereoVisionApp::ModularSBASolver::ProjectorModule::addProjectionCostFunction
//       libs/sparsesolver/modularsbasolver.h:169-177
//   PinholeCamProjModule::addProjectionCostFunction / init
//       libs/sparsesolver/sbamodules/pinholecameraprojectormodule.cpp:46-272, 302-533
//   SparseSolverBase step protocol (init / opt_step / writeResults / cleanup)
//       libs/sparsesolver/sparsesolverbase.cpp:71-88, modularsbasolver.cpp:679-951
//
// Parameter blocks are identified BY ADDRESS exactly as in the reference
// (pointer-stable InPlaceVector storage, modularsbasolver.h:925-935): the
// adaptor maps `double*` to dense indices and, at init(), packs the blocks
// into the flat arrays the C ABI takes; writeResults() scatters them back.
// No Eigen / Qt / Ceres types appear here; uv and stiffness are plain arrays.
#ifndef SBA_B200_ADAPTOR_HPP
#define SBA_B200_ADAPTOR_HPP

#include <array>
#include <cstdint>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "sba_b200.h"

namespace sba_b200 {

class GpuSBASolver;

// stands in for StereoVision::Geometry::RigidBodyTransform<double> (axis-angle r, translation t) in the signature below
struct RigidOffset {
    double r[3] = {0.0, 0.0, 0.0};
    double t[3] = {0.0, 0.0, 0.0};
    bool isIdentity() const {   // the reference's own test (pinholecameraprojectormodule.cpp:58): |r| < 1e-6 and |t| < 1e-6
        return r[0] * r[0] + r[1] * r[1] + r[2] * r[2] < 1e-12 && t[0] * t[0] + t[1] * t[1] + t[2] * t[2] < 1e-12;
    }
};

// Drop-in for PinholeCamProjModule: one instance per Camera datablock (= one
// camera model: f, pp, k, t, B blocks shared by all its images).
class GpuPinholeProjectorModule {
public:
    GpuPinholeProjectorModule(GpuSBASolver* solver, int32_t model) : _solver(solver), _model(model) {}

    // the full signature of ProjectorModule::addProjectionCostFunction (modularsbasolver.h:169-177), same argument
    // meaning; returns false when the module was not set up (pinholecameraprojectormodule.cpp:253-255), when a block
    // was not registered, or when the pose was registered under another camera model than this module's.
    // uv = pixel, stiffness = 2x2 row-major.
    inline bool addProjectionCostFunction(double* pointData, double* poseOrientation, double* posePosition,
                                          const double uv[2], const double stiffness[4],
                                          const RigidOffset& offset = RigidOffset(),
                                          double* leverArmOrientation = nullptr, double* leverArmPosition = nullptr);

    int32_t model() const { return _model; }
    // the module's intrinsic blocks after writeResults() (PinholeCamProjModule::writeResults reads the same five blocks back
    // into the Camera datablock, pinholecameraprojectormodule.cpp:535-600): f, pp[2], k[3], t[2], B[2]
    inline const double* intrinsics() const;

private:
    GpuSBASolver* _solver;
    int32_t _model;
};

class GpuSBASolver {
public:
    explicit GpuSBASolver(const sba_options* opt = nullptr) {
        if (opt) _opt = *opt; else sba_default_options(&_opt);
    }
    ~GpuSBASolver() { cleanup(); }
    GpuSBASolver(const GpuSBASolver&) = delete;
    GpuSBASolver& operator=(const GpuSBASolver&) = delete;

    // --- configuration: ModularSBASolver setters (modularsbasolver.h:876-884) ---
    void setOptimizationSteps(int n) { _opt.max_iterations = n; }
    void setFuncTolerance(double v) { _opt.function_tolerance = v; }
    void setParamsTolerance(double v) { _opt.parameter_tolerance = v; }
    // Qt property useSphericalCostPinholeCamProjCost (pinholecameraprojectormodule.cpp:28,257-271)
    void setUseSphericalCost(bool v) { _spherical = v; }
    const std::string& lastError() const { return _err; }

    // --- registration: getNodeForLandmark / getNodeForFrame (modularsbasolver.cpp:196-306) ---
    // pos / rAxis / t point into storage owned by the caller that must not move until cleanup().
    int32_t registerPoint(double* pos, bool fixed = false) {
        auto it = _ptIndex.find(pos);
        if (it != _ptIndex.end()) return it->second;
        const int32_t i = (int32_t)_ptPtr.size();
        _ptIndex[pos] = i; _ptPtr.push_back(pos); _ptFixed.push_back(fixed ? 1 : 0);
        return i;
    }
    int32_t registerPose(double* rAxis, double* t, int32_t cameraModel, bool fixed = false) {
        auto it = _poseIndex.find(rAxis);
        if (it != _poseIndex.end()) return it->second;
        const int32_t i = (int32_t)_poseR.size();
        _poseIndex[rAxis] = i; _poseR.push_back(rAxis); _poseT.push_back(t);
        _poseModel.push_back(cameraModel); _poseFixed.push_back(fixed ? 1 : 0);
        return i;
    }
    // PinholeCamProjModule::init intrinsic blocks; returns the module for this camera model
    // fixedMask: bit k set -> block k constant (k = 0 f, 1 pp, 2 k, 3 t, 4 B) -- Camera::isFixed / FixedParameter::CameraInternal /
    // use*DistortionModel of PinholeCamProjModule::init (pinholecameraprojectormodule.cpp:356-507); 0x1f = everything constant
    GpuPinholeProjectorModule addCameraModel(double f, const double pp[2], const double k[3], const double t[2],
                                             const double B[2], uint8_t fixedMask = 0x1f) {
        const double rec[10] = {f, pp[0], pp[1], k[0], k[1], k[2], t[0], t[1], B[0], B[1]};
        _intr.insert(_intr.end(), rec, rec + 10);
        _intrMask.push_back(fixedMask);
        return GpuPinholeProjectorModule(this, (int32_t)(_intr.size() / 10 - 1));
    }
    // FixedSizeNormalPrior<n,n> on an intrinsic block of a model (block 0 f, 1 pp, 2 k, 3 t, 4 B; A = dim x dim row-major)
    void addIntrinsicsPrior(int32_t model, int32_t block, const double* A, const double* b) {
        const int dim = block == 0 ? 1 : (block == 2 ? 3 : 2);
        double A9[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, b3[3] = {0, 0, 0};
        for (int i = 0; i < dim * dim; i++) A9[i] = A[i];
        for (int i = 0; i < dim; i++) b3[i] = b[i];
        _iprModel.push_back(model); _iprBlock.push_back(block);
        _iprA.insert(_iprA.end(), A9, A9 + 9); _iprB.insert(_iprB.end(), b3, b3 + 3);
    }
    // a mounting = the pair of lever-arm blocks (boresight, leverArm), identified by address like every block
    int32_t registerLeverArm(double* orientation, double* position, bool fixed = false) {
        auto it = _mountIndex.find(orientation);
        if (it != _mountIndex.end()) return it->second;
        const int32_t i = (int32_t)_mountR.size();
        _mountIndex[orientation] = i; _mountR.push_back(orientation); _mountT.push_back(position); _mountFixed.push_back(fixed ? 1 : 0);
        return i;
    }
    // problem.AddResidualBlock(new FixedSizeNormalPrior<3,3>(A, b), nullptr, block)
    bool addNormalPrior(double* block, const double A[9], const double b[3]) {
        int32_t kind, idx;
        auto ip = _ptIndex.find(block);
        if (ip != _ptIndex.end()) { kind = SBA_PRIOR_POINT; idx = ip->second; }
        else {
            auto ir = _poseIndex.find(block);
            if (ir != _poseIndex.end()) { kind = SBA_PRIOR_CAM_R; idx = ir->second; }
            else {
                idx = -1;
                for (size_t i = 0; i < _poseT.size(); i++) if (_poseT[i] == block) { idx = (int32_t)i; break; }
                if (idx < 0) { _err = "addNormalPrior: unknown parameter block"; return false; }
                kind = SBA_PRIOR_CAM_T;
            }
        }
        _priorKind.push_back(kind); _priorIndex.push_back(idx);
        _priorA.insert(_priorA.end(), A, A + 9); _priorB.insert(_priorB.end(), b, b + 3);
        return true;
    }

    // --- SteppedProcess protocol ---
    bool init() {
        if (_handle) return true;
        if (sba_create(&_opt, &_handle) != SBA_OK) { _err = sba_create_error(); _handle = nullptr; return false; }
        const int32_t nc = (int32_t)_poseR.size(), np = (int32_t)_ptPtr.size();
        const int32_t nmt = (int32_t)_mountR.size();
        _r.resize(3 * (size_t)nc); _t.resize(3 * (size_t)nc); _x.resize(3 * (size_t)np);
        _mb.resize(3 * (size_t)nmt); _ml.resize(3 * (size_t)nmt);
        _poseOffset.resize(6 * (size_t)nc, 0.0); _poseMount.resize(nc, -1);
        gather();
        std::vector<uint8_t> kind(nc, _spherical ? SBA_COST_SPHERICAL : SBA_COST_PLANE);
        bool anyOffset = false;
        for (double v : _poseOffset) anyOffset = anyOffset || v != 0.0;
        bool ok = sba_set_intrinsics_blocks(_handle, (int32_t)(_intr.size() / 10), _intr.data(), _intrMask.data()) == SBA_OK
               && sba_add_intrinsics_priors(_handle, (int32_t)_iprModel.size(), _iprModel.data(), _iprBlock.data(), _iprA.data(), _iprB.data()) == SBA_OK
               && sba_set_poses(_handle, _r.data(), _t.data(), nc, _poseModel.data(), _poseFixed.data(), kind.data()) == SBA_OK
               && sba_set_sensor_offsets(_handle, anyOffset ? _poseOffset.data() : nullptr) == SBA_OK
               && sba_set_lever_arms(_handle, nmt, _mb.data(), _ml.data(), _mountFixed.data(), _poseMount.data()) == SBA_OK
               && sba_set_points(_handle, _x.data(), np, _ptFixed.data()) == SBA_OK
               && sba_add_observations(_handle, (int32_t)_obsCam.size(), _obsCam.data(), _obsPt.data(), _obsUV.data(),
                                       _hasInfo ? _obsInfo.data() : nullptr) == SBA_OK
               && sba_add_priors(_handle, (int32_t)_priorKind.size(), _priorKind.data(), _priorIndex.data(),
                                 _priorA.data(), _priorB.data()) == SBA_OK
               && sba_finalize(_handle) == SBA_OK;
        if (!ok) { _err = sba_last_error(_handle); cleanup(); }
        return ok;
    }
    // first call runs the whole LM, later calls return true (modularsbasolver.cpp:886-888)
    bool opt_step() {
        if (!_handle) return false;
        if (_notFirstStep) return true;
        gather();
        if (sba_solve(_handle, 1, &_summary) != SBA_OK) { _err = sba_last_error(_handle); return false; }
        _notFirstStep = true;
        return true;
    }
    // nodes -> caller blocks (the reference narrows to float when writing datablocks; the blocks stay double)
    bool writeResults() {
        if (!_handle) return false;
        for (size_t i = 0; i < _poseR.size(); i++) {
            std::memcpy(_poseR[i], &_r[3 * i], 3 * sizeof(double));
            std::memcpy(_poseT[i], &_t[3 * i], 3 * sizeof(double));
        }
        for (size_t i = 0; i < _ptPtr.size(); i++) std::memcpy(_ptPtr[i], &_x[3 * i], 3 * sizeof(double));
        for (size_t i = 0; i < _mountR.size(); i++) {
            std::memcpy(_mountR[i], &_mb[3 * i], 3 * sizeof(double));
            std::memcpy(_mountT[i], &_ml[3 * i], 3 * sizeof(double));
        }
        return true;
    }
    void cleanup() {
        if (_handle) { sba_destroy(_handle); _handle = nullptr; }
        _notFirstStep = false;
    }
    const sba_summary& summary() const { return _summary; }
    sba_solver* handle() { return _handle; }

private:
    friend class GpuPinholeProjectorModule;
    void gather() {
        for (size_t i = 0; i < _poseR.size(); i++) {
            std::memcpy(&_r[3 * i], _poseR[i], 3 * sizeof(double));
            std::memcpy(&_t[3 * i], _poseT[i], 3 * sizeof(double));
        }
        for (size_t i = 0; i < _ptPtr.size(); i++) std::memcpy(&_x[3 * i], _ptPtr[i], 3 * sizeof(double));
        for (size_t i = 0; i < _mountR.size(); i++) {
            std::memcpy(&_mb[3 * i], _mountR[i], 3 * sizeof(double));
            std::memcpy(&_ml[3 * i], _mountT[i], 3 * sizeof(double));
        }
    }
    // the decorators of the observation (pinholecameraprojectormodule.cpp:58-214).  The engine keeps ONE offset and ONE
    // mounting per pose (what steviapp's image / trajectory modules pass: the offset and lever arm belong to the image), so a
    // second observation of the same pose with other decorators is refused.
    bool setPoseDecorators(int32_t pose, const RigidOffset& offset, double* leverR, double* leverT) {
        if (_poseOffset.size() < 6 * _poseR.size()) { _poseOffset.resize(6 * _poseR.size(), 0.0); _poseMount.resize(_poseR.size(), -1); _poseSeen.resize(_poseR.size(), 0); }
        double o6[6] = {0, 0, 0, 0, 0, 0};
        if (!offset.isIdentity()) { for (int a = 0; a < 3; a++) { o6[a] = offset.r[a]; o6[3 + a] = offset.t[a]; } }
        int32_t mount = -1;
        if (leverR != nullptr && leverT != nullptr) {
            auto it = _mountIndex.find(leverR);
            if (it == _mountIndex.end() || _mountT[it->second] != leverT) { _err = "addProjectionCostFunction: lever-arm blocks were not registered"; return false; }
            mount = it->second;
        }
        if (_poseSeen[pose]) {
            bool same = _poseMount[pose] == mount;
            for (int a = 0; a < 6; a++) same = same && _poseOffset[6 * (size_t)pose + a] == o6[a];
            if (!same) { _err = "addProjectionCostFunction: a pose must keep one sensor offset and one lever arm over its observations"; return false; }
            return true;
        }
        _poseSeen[pose] = 1; _poseMount[pose] = mount;
        for (int a = 0; a < 6; a++) _poseOffset[6 * (size_t)pose + a] = o6[a];
        return true;
    }
    bool addObservation(int32_t model, double* pt, double* rAxis, double* t, const double uv[2], const double W[4],
                        const RigidOffset& offset, double* leverR, double* leverT) {
        auto ip = _ptIndex.find(pt);
        auto ic = _poseIndex.find(rAxis);
        if (ip == _ptIndex.end() || ic == _poseIndex.end() || _poseT[ic->second] != t) {
            _err = "addProjectionCostFunction: parameter block was not registered";
            return false;
        }
        if (_poseModel[ic->second] != model) {   // in the reference the module OWNS the intrinsic blocks of its Camera
            _err = "addProjectionCostFunction: the pose was registered under another camera model than this module's";
            return false;
        }
        if (!setPoseDecorators(ic->second, offset, leverR, leverT)) return false;
        const bool identity = !W || (W[0] == 1.0 && W[1] == 0.0 && W[2] == 0.0 && W[3] == 1.0);
        if (!identity && !_hasInfo) {
            _obsInfo.assign(4 * _obsCam.size(), 0.0);
            for (size_t i = 0; i < _obsCam.size(); i++) { _obsInfo[4 * i] = 1.0; _obsInfo[4 * i + 3] = 1.0; }
            _hasInfo = true;
        }
        _obsCam.push_back(ic->second); _obsPt.push_back(ip->second);
        _obsUV.push_back(uv[0]); _obsUV.push_back(uv[1]);
        if (_hasInfo) {
            const double I[4] = {1, 0, 0, 1};
            const double* w = W ? W : I;
            _obsInfo.insert(_obsInfo.end(), w, w + 4);
        }
        return true;
    }

    sba_options _opt;
    sba_solver* _handle = nullptr;
    sba_summary _summary{};
    bool _spherical = false, _notFirstStep = false, _hasInfo = false;
    std::string _err;
    std::unordered_map<double*, int32_t> _ptIndex, _poseIndex, _mountIndex;
    std::vector<double*> _ptPtr, _poseR, _poseT, _mountR, _mountT;
    std::vector<uint8_t> _ptFixed, _poseFixed, _intrMask, _mountFixed, _poseSeen;
    std::vector<int32_t> _poseModel, _obsCam, _obsPt, _priorKind, _priorIndex, _poseMount, _iprModel, _iprBlock;
    std::vector<double> _intr, _obsUV, _obsInfo, _priorA, _priorB, _r, _t, _x, _poseOffset, _mb, _ml, _iprA, _iprB;
};

inline bool GpuPinholeProjectorModule::addProjectionCostFunction(double* pointData, double* poseOrientation,
                                                                 double* posePosition, const double uv[2],
                                                                 const double stiffness[4], const RigidOffset& offset,
                                                                 double* leverArmOrientation, double* leverArmPosition) {
    if (!_solver) return false;
    if ((leverArmOrientation == nullptr) != (leverArmPosition == nullptr)) {   // the reference takes the lever arm only when both are set
        leverArmOrientation = nullptr; leverArmPosition = nullptr;
    }
    // UVCost / LeverArmCost / PoseTransformCost / PoseTransformLeverArmCost (pinholecameraprojectormodule.cpp:58-214)
    return _solver->addObservation(_model, pointData, poseOrientation, posePosition, uv, stiffness, offset, leverArmOrientation, leverArmPosition);
}
inline const double* GpuPinholeProjectorModule::intrinsics() const { return _solver ? &_solver->_intr[10 * (size_t)_model] : nullptr; }

}  // namespace sba_b200
#endif
