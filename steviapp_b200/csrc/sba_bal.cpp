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
#include <new>
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

extern "C" {

const char* sba_bal_error(void) { return g_bal_error.c_str(); }

int sba_bal_open(const char* path, sba_bal** out) {
    if (!path || !out) return bal_fail("sba_bal_open: null argument");
    *out = nullptr;
    std::vector<char> buf;
    if (!slurp(path, buf)) return bal_fail(std::string("sba_bal_open: cannot read ") + path);
    Cursor c{buf.data()};
    const long nc = c.integer(), np = c.integer(), no = c.integer();
    if (!c.ok || nc <= 0 || np <= 0 || no < 0 || nc > 0x7fffffffL || np > 0x7fffffffL || no > 0x7fffffffL)
        return bal_fail("sba_bal_open: bad header (expected: cameras points observations)");
    sba_bal* b = new (std::nothrow) sba_bal;
    if (!b) return bal_fail("sba_bal_open: out of memory");
    b->n_cams = (int32_t)nc; b->n_pts = (int32_t)np; b->n_obs = (int32_t)no;
    b->cam.resize((size_t)no); b->pt.resize((size_t)no); b->xy.resize(2 * (size_t)no);
    b->cam9.resize(9 * (size_t)nc); b->xyz.resize(3 * (size_t)np);
    for (long i = 0; i < no && c.ok; i++) {
        const long ci = c.integer(), pi = c.integer();
        b->xy[2 * (size_t)i] = c.real(); b->xy[2 * (size_t)i + 1] = c.real();
        if (c.ok && (ci < 0 || ci >= nc || pi < 0 || pi >= np)) {   // the reference asserts the same bounds (bal_example.cpp:312-316)
            delete b;
            return bal_fail("sba_bal_open: observation " + std::to_string(i) + " has an index out of bounds");
        }
        b->cam[(size_t)i] = (int32_t)ci; b->pt[(size_t)i] = (int32_t)pi;
    }
    for (size_t k = 0; k < b->cam9.size() && c.ok; k++) b->cam9[k] = c.real();
    for (size_t k = 0; k < b->xyz.size() && c.ok; k++) b->xyz[k] = c.real();
    if (!c.ok) { delete b; return bal_fail("sba_bal_open: file ends early or holds a token that is not a number"); }
    *out = b;
    return SBA_OK;
}

void sba_bal_close(sba_bal* b) { delete b; }

int sba_bal_sizes(const sba_bal* b, int32_t* n_cams, int32_t* n_pts, int32_t* n_obs) {
    if (!b) return bal_fail("sba_bal_sizes: null handle");
    if (n_cams) *n_cams = b->n_cams;
    if (n_pts) *n_pts = b->n_pts;
    if (n_obs) *n_obs = b->n_obs;
    return SBA_OK;
}

int sba_bal_raw(const sba_bal* b, int32_t* cam_idx, int32_t* pt_idx, double* xy, double* cam9, double* xyz) {
    if (!b) return bal_fail("sba_bal_raw: null handle");
    if (cam_idx) std::memcpy(cam_idx, b->cam.data(), b->cam.size() * sizeof(int32_t));
    if (pt_idx) std::memcpy(pt_idx, b->pt.data(), b->pt.size() * sizeof(int32_t));
    if (xy) std::memcpy(xy, b->xy.data(), b->xy.size() * sizeof(double));
    if (cam9) std::memcpy(cam9, b->cam9.data(), b->cam9.size() * sizeof(double));
    if (xyz) std::memcpy(xyz, b->xyz.data(), b->xyz.size() * sizeof(double));
    return SBA_OK;
}

int sba_bal_convert(const sba_bal* b, double* raxis, double* t, double* intr10, int32_t* cam_model, double* xyz,
                    int32_t* cam_idx, int32_t* pt_idx, double* uv, int32_t* n_undistort_failed) {
    if (!b || !raxis || !t || !intr10 || !cam_model || !xyz || !cam_idx || !pt_idx || !uv) return bal_fail("sba_bal_convert: null argument");
    for (int32_t i = 0; i < b->n_cams; i++) {
        const double* c9 = &b->cam9[9 * (size_t)i];
        double R[9], Re[9];
        rodrigues(c9, R);
        // R_e = R^T F, F = diag(1, -1, -1): column j of R^T scaled by F_jj
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) Re[3 * r + c] = R[3 * c + r] * (c == 0 ? 1.0 : -1.0);
        log_rotation(Re, raxis + 3 * (size_t)i);
        for (int r = 0; r < 3; r++) t[3 * (size_t)i + r] = -(R[r] * c9[3] + R[3 + r] * c9[4] + R[6 + r] * c9[5]);   // -R^T t
        double* m = intr10 + (size_t)SBA_INTR_STRIDE * i;
        for (int k = 0; k < SBA_INTR_STRIDE; k++) m[k] = 0.0;
        m[0] = c9[6];
        cam_model[i] = i;
    }
    std::memcpy(xyz, b->xyz.data(), b->xyz.size() * sizeof(double));
    int32_t failed = 0;
    for (int32_t i = 0; i < b->n_obs; i++) {
        const double* c9 = &b->cam9[9 * (size_t)b->cam[(size_t)i]];
        const double f = c9[6], k1 = c9[7], k2 = c9[8];
        const double xd = b->xy[2 * (size_t)i] / f, yd = b->xy[2 * (size_t)i + 1] / f;   // distorted normalised coordinates
        const double rd = std::sqrt(xd * xd + yd * yd);
        // radius before distortion: rho (1 + k1 rho^2 + k2 rho^4) = rd, Newton from rho = rd
        double rho = rd;
        bool ok = true;
        for (int it = 0; it < 50; it++) {
            const double r2 = rho * rho, g = rho * (1.0 + r2 * (k1 + k2 * r2)) - rd, dg = 1.0 + r2 * (3.0 * k1 + 5.0 * k2 * r2);
            if (!(std::fabs(dg) > 1e-12)) { ok = false; break; }
            const double step = g / dg;
            rho -= step;
            if (std::fabs(step) <= 1e-15 * std::fmax(rho, 1.0)) break;
        }
        // accepted when the radius equation holds (a wide-angle observation beyond the turning point of the polynomial has no
        // root near rd: kept distorted and counted)
        if (ok) { const double r2 = rho * rho; ok = std::fabs(rho * (1.0 + r2 * (k1 + k2 * r2)) - rd) <= 1e-12 * std::fmax(rd, 1.0); }
        if (!ok || !(rho >= 0.0) || !std::isfinite(rho)) { rho = rd; failed++; }
        const double s = rd > 0.0 ? rho / rd : 1.0;
        uv[2 * (size_t)i] = f * xd * s;
        uv[2 * (size_t)i + 1] = -f * yd * s;
        cam_idx[i] = b->cam[(size_t)i]; pt_idx[i] = b->pt[(size_t)i];
    }
    if (n_undistort_failed) *n_undistort_failed = failed;
    return SBA_OK;
}

// The file's OWN objective (bal_example.cpp:168-204): 1/2 sum |f (1 + k1 |p|^2 + k2 |p|^4) p - xy|^2, p = -P.xy / P.z,
// P = R(r) X + t -- so that costs before / after a solve can be quoted in BAL's metric (the engine minimises the residual
// with the distortion moved to the measurement, which is only first-order equivalent).  cam9 / xyz: NULL = the file's.
int sba_bal_cost(const sba_bal* b, const double* cam9, const double* xyz, double* cost, double* rms_px) {
    if (!b) return bal_fail("sba_bal_cost: null handle");
    if (!cam9) cam9 = b->cam9.data();
    if (!xyz) xyz = b->xyz.data();
    double sum = 0.0;
    for (int32_t i = 0; i < b->n_obs; i++) {
        const double* c9 = cam9 + 9 * (size_t)b->cam[(size_t)i];
        const double* X = xyz + 3 * (size_t)b->pt[(size_t)i];
        double R[9];
        rodrigues(c9, R);
        const double Px = R[0] * X[0] + R[1] * X[1] + R[2] * X[2] + c9[3], Py = R[3] * X[0] + R[4] * X[1] + R[5] * X[2] + c9[4],
                     Pz = R[6] * X[0] + R[7] * X[1] + R[8] * X[2] + c9[5];
        const double px = -Px / Pz, py = -Py / Pz, r2 = px * px + py * py;
        const double s = c9[6] * (1.0 + r2 * (c9[7] + c9[8] * r2));
        const double ex = s * px - b->xy[2 * (size_t)i], ey = s * py - b->xy[2 * (size_t)i + 1];
        sum += ex * ex + ey * ey;
    }
    if (cost) *cost = 0.5 * sum;
    if (rms_px) *rms_px = b->n_obs ? std::sqrt(sum / (2.0 * b->n_obs)) : 0.0;
    return SBA_OK;
}

// engine-convention poses (sba_bal_convert's raxis / t after a solve) back to BAL cameras: R = F R_e^T, t = -R t_e with
// F = diag(1, -1, -1); f, k1, k2 are the file's (held constant by the conversion)
int sba_bal_cameras_from_engine(const sba_bal* b, const double* raxis, const double* t, double* cam9_out) {
    if (!b || !raxis || !t || !cam9_out) return bal_fail("sba_bal_cameras_from_engine: null argument");
    for (int32_t i = 0; i < b->n_cams; i++) {
        double Re[9], R[9];
        rodrigues(raxis + 3 * (size_t)i, Re);
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) R[3 * r + c] = Re[3 * c + r] * (r == 0 ? 1.0 : -1.0);   // F R_e^T
        double* o = cam9_out + 9 * (size_t)i;
        log_rotation(R, o);
        const double* te = t + 3 * (size_t)i;
        for (int r = 0; r < 3; r++) o[3 + r] = -(R[3 * r] * te[0] + R[3 * r + 1] * te[1] + R[3 * r + 2] * te[2]);
        o[6] = b->cam9[9 * (size_t)i + 6]; o[7] = b->cam9[9 * (size_t)i + 7]; o[8] = b->cam9[9 * (size_t)i + 8];
    }
    return SBA_OK;
}

int sba_bal_write(const char* path, int32_t n_cams, int32_t n_pts, int32_t n_obs, const int32_t* cam_idx, const int32_t* pt_idx,
                  const double* xy, const double* cam9, const double* xyz) {
    if (!path || !cam_idx || !pt_idx || !xy || !cam9 || !xyz || n_cams <= 0 || n_pts <= 0 || n_obs < 0) return bal_fail("sba_bal_write: bad argument");
    FILE* f = std::fopen(path, "w");
    if (!f) return bal_fail(std::string("sba_bal_write: cannot open ") + path);
    std::fprintf(f, "%d %d %d\n", n_cams, n_pts, n_obs);
    for (int32_t i = 0; i < n_obs; i++)
        std::fprintf(f, "%d %d %.17g %.17g\n", cam_idx[i], pt_idx[i], xy[2 * (size_t)i], xy[2 * (size_t)i + 1]);
    for (size_t k = 0; k < 9 * (size_t)n_cams; k++) std::fprintf(f, "%.17g\n", cam9[k]);
    for (size_t k = 0; k < 3 * (size_t)n_pts; k++) std::fprintf(f, "%.17g\n", xyz[k]);
    const bool bad = std::ferror(f) != 0;
    if (std::fclose(f) != 0 || bad) return bal_fail(std::string("sba_bal_write: write failed on ") + path);
    return SBA_OK;
}

}  // extern "C"
