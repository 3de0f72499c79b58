// hostsim.cpp -- TEST-ONLY host build of the per-thread device functions (csrc/*.cuh compiled
// with g++, no CUDA). Lets the CPU-only test suite check the kernels' arithmetic against the
// oracle before any GPU time is spent. Never linked into libsfm_b200.so.
#include "../../structure-from-motion-_b200/csrc/f_math.cuh"
#include "../../structure-from-motion-_b200/csrc/hostsim_extra.inc"

using namespace sfm;

extern "C" void hs_f8pt(const double *x1x2 /*[8,4]*/, double *F) {
    double px1[8], py1[8], px2[8], py2[8];
    for (int i = 0; i < 8; ++i) {
        px1[i] = x1x2[4 * i]; py1[i] = x1x2[4 * i + 1]; px2[i] = x1x2[4 * i + 2]; py2[i] = x1x2[4 * i + 3];
    }
    f8pt_from_sample(px1, py1, px2, py2, F);
}

extern "C" void hs_f_score(const double *x1x2, int n, const double *F, double thr, unsigned char *mask,
                           double *err) {
    FMat f = {F[0], F[1], F[2], F[3], F[4], F[5], F[6], F[7], F[8]};
    for (int i = 0; i < n; ++i) {
        const double *r = x1x2 + 4 * i;
        mask[i] = f_is_inlier(f, r[0], r[1], r[2], r[3], 2.0 * thr) ? 1 : 0;
        err[i] = f_error_exact(f, r[0], r[1], r[2], r[3]);
    }
}
