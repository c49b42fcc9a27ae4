// sba_bal.cpp -- reader / writer of BAL ("Bundle Adjustment in the Large") text problems and their conversion to the
// engine's conventions (the loader that lets real Ladybug / Venice files replace the synthetic shapes).
// Host code only, no CUDA.  The file format and the camera model are the ones of the reference's in-tree BAL example
// (modules/g2o/g2o/examples/bal/bal_example.cpp:271-340 for the format, :168-204 for the projection):
//     header   n_cams n_pts n_obs
//     n_obs  x cam_index point_index x y
//     n_cams x 9 doubles: Rodrigues r[3] (world -> camera), t[3], f, k1, k2
//     n_pts  x 3 doubles
//     P = R(r) X + t;  p = -P.xy / P.z;  pixel = f (1 + k1 |p|^2 + k2 |p|^4) p         (camera looks down -z)
//
// Conversion to the steviapp conventions the engine keeps (include/sba_b200.h): poses are sensor -> world, the camera looks
// down +z, and the lens model is applied to the MEASURED pixel (sbamodules/pinholecameraprojectormodule.h:28-71).  With
// F = diag(1, -1, -1):  Pbar = F P  =>  R_e = R(r)^T F,  t_e = -R(r)^T t (the camera centre), and a measured (x, y) becomes
// (u, v) = (x_u, -y_u) where (x_u, y_u) = f p is the pixel with the file's radial distortion removed (Newton on the
// radius; f, k1, k2 are held constant -- FixedParameter::CameraInternal -- so this happens once at load time).  Every camera
// gets its own intrinsics model {f, pp = 0, k = 0, t = 0, B = 0}.  The plane-cost residual of the engine,
// f Pbar.xy / Pbar.z - (u, v), then equals the BAL residual divided by the local distortion factor: the same minimiser for
// exact data, first-order equivalent otherwise (documented in DESIGN.md).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "sba_b200.h"

struct sba_bal {
    int32_t n_cams = 0, n_pts = 0, n_obs = 0;
    std::vector<int32_t> cam, pt;
    std::vector<double> xy, cam9, xyz;
};

static thread_local std::string g_bal_error;
static int bal_fail(const std::string& m) { g_bal_error = m; return SBA_ERR_INVALID; }

namespace {

// whole file in memory, then strtod / strtol over it: a Venice-size file (5 M observations) parses in about a second
bool slurp(const char* path, std::vector<char>& buf) {
    FILE* f = std::fopen(path, "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (n < 0) { std::fclose(f); return false; }
    buf.resize((size_t)n + 1);
    const size_t got = std::fread(buf.data(), 1, (size_t)n, f);
    std::fclose(f);
    buf[got] = 0;
    buf.resize(got + 1);
    return true;
}

struct Cursor {
    const char* p;
    bool ok = true;
    long integer() {
        char* e = nullptr;
        const long v = std::strtol(p, &e, 10);
        if (e == p) ok = false;
        p = e;
        return v;
    }
    double real() {
        char* e = nullptr;
        const double v = std::strtod(p, &e);
        if (e == p) ok = false;
        p = e;
        return v;
    }
};

void rodrigues(const double r[3], double R[9]) {   // same form as bal_example.cpp:174-190 (Taylor branch at theta = 0)
    const double th2 = r[0] * r[0] + r[1] * r[1] + r[2] * r[2];
    double a, b;   // R = I + a [r]x + b [r]x^2
    if (th2 > 1e-16) { const double th = std::sqrt(th2); a = std::sin(th) / th; b = (1.0 - std::cos(th)) / th2; }
    else { a = 1.0 - th2 / 6.0; b = 0.5 - th2 / 24.0; }
    const double x = r[0], y = r[1], z = r[2];
    R[0] = 1.0 - b * (y * y + z * z); R[1] = -a * z + b * x * y;        R[2] = a * y + b * x * z;
    R[3] = a * z + b * x * y;         R[4] = 1.0 - b * (x * x + z * z); R[5] = -a * x + b * y * z;
    R[6] = -a * y + b * x * z;        R[7] = a * x + b * y * z;         R[8] = 1.0 - b * (x * x + y * y);
}

// rotation matrix -> axis-angle through the unit quaternion (largest-component branch): stable near pi, where R^T F lands
// whenever the BAL rotation is small
void log_rotation(const double R[9], double r[3]) {
    double q[4];   // w, x, y, z
    const double tr = R[0] + R[4] + R[8];
    if (tr > 0.0) {
        const double s = 2.0 * std::sqrt(tr + 1.0);
        q[0] = 0.25 * s; q[1] = (R[7] - R[5]) / s; q[2] = (R[2] - R[6]) / s; q[3] = (R[3] - R[1]) / s;
    } else if (R[0] > R[4] && R[0] > R[8]) {
        const double s = 2.0 * std::sqrt(1.0 + R[0] - R[4] - R[8]);
        q[0] = (R[7] - R[5]) / s; q[1] = 0.25 * s; q[2] = (R[1] + R[3]) / s; q[3] = (R[2] + R[6]) / s;
    } else if (R[4] > R[8]) {
        const double s = 2.0 * std::sqrt(1.0 + R[4] - R[0] - R[8]);
        q[0] = (R[2] - R[6]) / s; q[1] = (R[1] + R[3]) / s; q[2] = 0.25 * s; q[3] = (R[5] + R[7]) / s;
    } else {
        const double s = 2.0 * std::sqrt(1.0 + R[8] - R[0] - R[4]);
        q[0] = (R[3] - R[1]) / s; q[1] = (R[2] + R[6]) / s; q[2] = (R[5] + R[7]) / s; q[3] = 0.25 * s;
    }
    if (q[0] < 0.0) for (double& v : q) v = -v;
    const double n = std::sqrt(q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    const double k = n > 1e-12 ? 2.0 * std::atan2(n, q[0]) / n : 2.0;
    r[0] = k * q[1]; r[1] = k * q[2]; r[2] = k * q[3];
}

}  // namespace
