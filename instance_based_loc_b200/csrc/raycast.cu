// raycast.cu -- synthetic RGB-D frames (SURVEY.md section 8(d) inputs, 8(f) row S1): one ray per pixel against the room's
// six planes and m oriented boxes / ellipsoids.  This replaces the reference's habitat-sim rendering
// (hm3d_data_generation.py; no graphics stack on this box) as plain compute; it only PRODUCES benchmark / test inputs
// and is never inside a timed region.  The arithmetic restates instance_based_loc_b200/synth.py:render (torch tensor
// ops, f64) step by step; the two agree to rounding (dot products there go through BLAS, whose summation order is not
// specified), which is what tests/test_z_memory_build_gpu.py::test_raycast_kernel_vs_tensor_ops checks.
//
// Pixel (r, c) has the camera-frame direction (hor[c] / f, ver[r] / f, 1) with hor = linspace(-W/2, W/2, W) and
// ver = linspace(H/2, -H/2, H) as FLOAT32 values (utils/depth_utils.py:57-63 back-projects with exactly these), NOT
// normalised, so the ray parameter of a hit is the camera depth Z.  Thread per pixel; the scene (<= a few hundred
// objects, 128 B each) is read through the read-only path with the whole warp on the same address.
#include "ibl_common.cuh"

struct RaycastCam {
    double R[9];   // camera -> world rotation, row-major
    double o[3];   // camera centre in the world
};

__device__ __forceinline__ double rc_dot(const double *a, double x, double y, double z) {
    return __dadd_rn(__dadd_rn(__dmul_rn(a[0], x), __dmul_rn(a[1], y)), __dmul_rn(a[2], z));
}

__global__ void __launch_bounds__(256) raycast_kernel(RaycastCam cam, int H, int W, double f, double lin_w, double lin_h,
                                                      const int32_t *__restrict__ kind, const double *__restrict__ half,
                                                      const double *__restrict__ Ro, const double *__restrict__ to, int m,
                                                      double room, double height, double *__restrict__ out_t,
                                                      int32_t *__restrict__ out_inst, double *__restrict__ out_nrm) {
    const int64_t P = (int64_t)H * W;
    for (int64_t pix = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; pix < P; pix += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(pix / W), c = (int)(pix % W);
        // numpy.linspace(start, stop, n): start + i * step in f64, cast to f32 (the last sample is set to stop exactly when n > 1)
        const double hx = (W > 1 && c == W - 1) ? 0.5 * W : (double)(float)(-0.5 * W + c * lin_w);
        const double vy = (H > 1 && r == H - 1) ? -0.5 * H : (double)(float)(0.5 * H - r * lin_h);
        const double dcx = hx / f, dcy = vy / f;
        const double dx = rc_dot(cam.R + 0, dcx, dcy, 1.0), dy = rc_dot(cam.R + 3, dcx, dcy, 1.0),
                     dz = rc_dot(cam.R + 6, dcx, dcy, 1.0);
        double best = INFINITY, nx = 0.0, ny = 0.0, nz = 0.0;
        int inst = -1;
        // room: floor z = 0, ceiling, four walls -- (normal, offset) with the plane n.x = offset, hit from the front only
        const double pn[6][3] = {{0, 0, 1}, {0, 0, -1}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}};
        const double po[6] = {0.0, -height, -room, -room, -room, -room};
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const double den = rc_dot(pn[k], dx, dy, dz);
            const double tt = (po[k] - rc_dot(pn[k], cam.o[0], cam.o[1], cam.o[2])) / den;
            if (den < 0.0 && tt > 1e-6 && tt < best) { best = tt; nx = pn[k][0]; ny = pn[k][1]; nz = pn[k][2]; }
        }
        for (int j = 0; j < m; ++j) {
            const double *R = Ro + 9 * j;                       // object -> world, row-major: local = R^T (x - t)
            const double hx_ = half[3 * j], hy_ = half[3 * j + 1], hz_ = half[3 * j + 2];
            const double ex = cam.o[0] - to[3 * j], ey = cam.o[1] - to[3 * j + 1], ez = cam.o[2] - to[3 * j + 2];
            const double ol[3] = {__dadd_rn(__dadd_rn(__dmul_rn(ex, R[0]), __dmul_rn(ey, R[3])), __dmul_rn(ez, R[6])),
                                  __dadd_rn(__dadd_rn(__dmul_rn(ex, R[1]), __dmul_rn(ey, R[4])), __dmul_rn(ez, R[7])),
                                  __dadd_rn(__dadd_rn(__dmul_rn(ex, R[2]), __dmul_rn(ey, R[5])), __dmul_rn(ez, R[8]))};
            const double dl[3] = {__dadd_rn(__dadd_rn(__dmul_rn(dx, R[0]), __dmul_rn(dy, R[3])), __dmul_rn(dz, R[6])),
                                  __dadd_rn(__dadd_rn(__dmul_rn(dx, R[1]), __dmul_rn(dy, R[4])), __dmul_rn(dz, R[7])),
                                  __dadd_rn(__dadd_rn(__dmul_rn(dx, R[2]), __dmul_rn(dy, R[5])), __dmul_rn(dz, R[8]))};
            const double hf[3] = {hx_, hy_, hz_};
            double tt, nl[3] = {0.0, 0.0, 0.0};
            bool ok;
            if (kind[j] == 0) {                                 // box: slab method
                double tnear = -INFINITY, tfar = INFINITY;
                int ax = 0;
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const double inv = 1.0 / dl[a];
                    const double t0 = (-hf[a] - ol[a]) * inv, t1 = (hf[a] - ol[a]) * inv;
                    const double tn = fmin(t0, t1), tf = fmax(t0, t1);
                    if (tn > tnear) { tnear = tn; ax = a; }     // first axis wins a tie (torch.max returns the first index)
                    tfar = fmin(tfar, tf);
                }
                ok = tnear < tfar && tnear > 1e-6 && tnear < best;
                const double s = dl[ax] > 0.0 ? 1.0 : (dl[ax] < 0.0 ? -1.0 : 0.0);
                nl[ax] = -s;
                tt = tnear;
            } else {                                            // ellipsoid: unit sphere in scaled coordinates
                const double os0 = ol[0] / hf[0], os1 = ol[1] / hf[1], os2 = ol[2] / hf[2];
                const double ds0 = dl[0] / hf[0], ds1 = dl[1] / hf[1], ds2 = dl[2] / hf[2];
                const double a = __dadd_rn(__dadd_rn(__dmul_rn(ds0, ds0), __dmul_rn(ds1, ds1)), __dmul_rn(ds2, ds2));
                const double b = __dadd_rn(__dadd_rn(__dmul_rn(os0, ds0), __dmul_rn(os1, ds1)), __dmul_rn(os2, ds2));
                const double cc = __dadd_rn(__dadd_rn(__dmul_rn(os0, os0), __dmul_rn(os1, os1)), __dmul_rn(os2, os2)) - 1.0;
                const double disc = __dsub_rn(__dmul_rn(b, b), __dmul_rn(a, cc));
                tt = (-b - sqrt(fmax(disc, 0.0))) / a;
                ok = disc > 0.0 && tt > 1e-6 && tt < best;
                double h0 = __dadd_rn(ol[0], __dmul_rn(tt, dl[0])), h1 = __dadd_rn(ol[1], __dmul_rn(tt, dl[1])),
                       h2 = __dadd_rn(ol[2], __dmul_rn(tt, dl[2]));
                h0 /= __dmul_rn(hf[0], hf[0]); h1 /= __dmul_rn(hf[1], hf[1]); h2 /= __dmul_rn(hf[2], hf[2]);
                const double nn = fmax(sqrt(__dadd_rn(__dadd_rn(__dmul_rn(h0, h0), __dmul_rn(h1, h1)), __dmul_rn(h2, h2))), 1e-12);
                nl[0] = h0 / nn; nl[1] = h1 / nn; nl[2] = h2 / nn;
            }
            if (ok) {
                best = tt;
                inst = j;
                nx = rc_dot(R + 0, nl[0], nl[1], nl[2]);        // local -> world: R n
                ny = rc_dot(R + 3, nl[0], nl[1], nl[2]);
                nz = rc_dot(R + 6, nl[0], nl[1], nl[2]);
            }
        }
        out_t[pix] = best;
        out_inst[pix] = inst;
        out_nrm[3 * pix] = nx; out_nrm[3 * pix + 1] = ny; out_nrm[3 * pix + 2] = nz;
    }
}

IBL_API int ibl_raycast_scene(const double *h_Twc16, int32_t H, int32_t W, double f, const int32_t *kind, const double *half,
                              const double *R, const double *t, int32_t m, double room, double height, double *out_t,
                              int32_t *out_inst, double *out_nrm, void *stream) {
    IBL_ARG(h_Twc16 != nullptr && H >= 1 && W >= 1 && f > 0 && m >= 0 && out_t && out_inst && out_nrm);
    IBL_ARG(m == 0 || (kind && half && R && t));
    RaycastCam cam;
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) cam.R[3 * i + j] = h_Twc16[4 * i + j];
        cam.o[i] = h_Twc16[4 * i + 3];
    }
    const int64_t P = (int64_t)H * W;
    int nb = (int)((P + 255) / 256);
    if (nb > IBL_SMS * 16) nb = IBL_SMS * 16;
    // numpy.linspace step: (stop - start) / (n - 1) in f64
    const double lin_w = W > 1 ? (double)W / (double)(W - 1) : 0.0, lin_h = H > 1 ? (double)H / (double)(H - 1) : 0.0;
    IBL_PROF(IBL_PROF_MISC, ibl_stream(stream), 44.0 * P);
    raycast_kernel<<<nb, 256, 0, ibl_stream(stream)>>>(cam, H, W, f, lin_w, lin_h, kind, half, R, t, m, room, height, out_t,
                                                       out_inst, out_nrm);
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}
