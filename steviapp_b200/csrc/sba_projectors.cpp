// sba_projectors.cpp -- host-side conversions that let other UV projectors of the reference run on the pinhole kernels when
// their intrinsics are held constant (FixedParameter::CameraInternal): the cost UV2ParametrizedXYZCost only sees the
// projector through the view ray d = dirFromUV(uv; params) (costfunctors/modularuvprojection.h:389-391), and the pinhole
// record {f, pp = 0, k = 0, t = 0, B = 0} maps a pixel (u, v) to d = (u, v, f).  Feeding (d_x, d_y) as the pixel therefore
// reproduces any projector whose third ray component is the focal length.  Host code only, no CUDA.
#include <cstdint>

#include "sba_b200.h"

extern "C" {

// PinholePushbroomUVProjector::dirFromUV (sbamodules/pinholecameraprojectormodule.h:79-116):
//   s = (u - W/2) / W;  du = sum a_k s^k, dv = sum b_k s^k (k = 0..5);  d = (u - pp + du, v + dv, f)
int sba_pushbroom_equivalent_uv(int32_t n, const double* uv, double sensor_width, double f, double pp, const double* as6,
                                const double* bs6, double* uv_out, double* intr10_out) {
    if (n < 0 || (n > 0 && (!uv || !uv_out)) || !as6 || !bs6 || !(sensor_width > 0.0) || !(f > 0.0)) return SBA_ERR_INVALID;
    for (int32_t i = 0; i < n; i++) {
        const double u = uv[2 * (size_t)i], v = uv[2 * (size_t)i + 1];
        const double s = (u - sensor_width / 2.0) / sensor_width;
        const double s2 = s * s, s3 = s2 * s, s4 = s3 * s, s5 = s4 * s;
        const double du = as6[0] + as6[1] * s + as6[2] * s2 + as6[3] * s3 + as6[4] * s4 + as6[5] * s5;
        const double dv = bs6[0] + bs6[1] * s + bs6[2] * s2 + bs6[3] * s3 + bs6[4] * s4 + bs6[5] * s5;
        uv_out[2 * (size_t)i] = u - pp + du;
        uv_out[2 * (size_t)i + 1] = v + dv;
    }
    if (intr10_out) {
        for (int k = 0; k < SBA_INTR_STRIDE; k++) intr10_out[k] = 0.0;
        intr10_out[0] = f;
    }
    return SBA_OK;
}

}  // extern "C"
