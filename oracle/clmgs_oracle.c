/*
 * clmgs_oracle.c -- CPU restatement of the CLM-GS per-camera-batch hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under clm-gs_b200/ may include, link or
 * call this file.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker / the
 * timed CPU baseline.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * /root/reference).  Shorthand: CSRC = submodules/gsplat/gsplat/cuda/csrc,
 * CK = submodules/clm_kernels, ENG = strategies/clm_offload/engine.py.
 *
 * Arithmetic contract (DESIGN.md "Arithmetic"): all decision-making float
 * math (cull masks, radii, tile rectangles, depth sort keys) is plain IEEE-754
 * binary32 with no FMA contraction (compile with -ffp-contract=off), evaluated
 * in the order glm expands the reference's expressions.  rsqrt() is 1/sqrt().
 * The reference itself is built with --use_fast_math (setup.py:81), so against
 * the *GPU* reference these stages can differ on boundary Gaussians; against
 * this oracle the CUDA path is bit-exact.  Parity pin: tests/test_oracle_pin.py
 * checks this file against gsplat's own _torch_impl.py twin (imported from the
 * reference) via fixtures in tests/golden/.  cpu-adam, fast-tsp and the
 * clm_kernels logic have no reference tests or golden vectors: those parts are
 * "parity unpinned"  and are pinned against torch.optim.Adam /
 * first principles instead.
 *
 * Build: make -C oracle   (gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------ */
/* Deterministic activations (shared op order with csrc/gs_math.cuh).        */
/* The reference uses torch.exp / torch.sigmoid / F.normalize                */
/* (strategies/base_gaussian_model.py:42-50); these are within 2 ulp of      */
/* them and identical on CPU and GPU by construction (only +,*,fma,rint).    */
/* ------------------------------------------------------------------------ */
ORC_API float orc_expf(float x) {
    if (x > 88.72283f) return INFINITY;
    if (x < -103.0f) return 0.0f;
    const float LOG2E = 1.442695040888963f;
    const float LN2_HI = 0.693145751953125f;      /* 12 high bits of ln2 */
    const float LN2_LO = 1.428606765330187e-06f;  /* ln2 - LN2_HI */
    float n = rintf(x * LOG2E);
    float r = fmaf(-n, LN2_HI, x);
    r = fmaf(-n, LN2_LO, r);
    /* degree-7 Taylor on |r| <= ln2/2 (truncation error < 6e-9 relative) */
    float p = 1.9841270e-4f;
    p = fmaf(p, r, 1.3888889e-3f);
    p = fmaf(p, r, 8.3333333e-3f);
    p = fmaf(p, r, 4.1666668e-2f);
    p = fmaf(p, r, 1.6666667e-1f);
    p = fmaf(p, r, 0.5f);
    p = fmaf(p, r, 1.0f);
    p = fmaf(p, r, 1.0f);
    /* scale by 2^n in two steps so that subnormal results stay correct */
    int ni = (int)n;
    int n1 = ni / 2, n2 = ni - n1;
    union { uint32_t u; float f; } s1, s2;
    s1.u = (uint32_t)(n1 + 127) << 23;
    s2.u = (uint32_t)(n2 + 127) << 23;
    return (p * s1.f) * s2.f;
}

ORC_API float orc_sigmoidf(float x) { return 1.0f / (1.0f + orc_expf(-x)); }

/* deterministic log shared with csrc/gs_math.cuh det_logf (stands in for torch.log in the
 * densification step: utils/general_utils.py:145 inverse_sigmoid, scaling_inverse_activation) */
ORC_API float orc_logf(float x) {
    if (!(x >= 0.0f)) return NAN;
    if (x == 0.0f) return -INFINITY;
    if (x == INFINITY) return x;
    int bias = 0;
    if (x < 1.17549435e-38f) { x *= 8388608.0f; bias = -23; }
    union { int32_t i; float f; } u, m;
    u.f = x;
    int32_t e = (u.i - 0x3f3504f3) >> 23;
    m.i = u.i - (e << 23);
    float f = m.f - 1.0f;
    float s = f / (2.0f + f);
    float z = s * s;
    float p = 0.22222222f;
    p = fmaf(p, z, 0.2857143f);
    p = fmaf(p, z, 0.4f);
    p = fmaf(p, z, 0.6666667f);
    float logm = fmaf(s * z, p, 2.0f * s);
    float fe = (float)(e + bias);
    return fmaf(fe, 0.693145751953125f, fmaf(fe, 1.428606765330187e-06f, logm));
}

ORC_API void orc_logf_array(int64_t n, const float *x, float *y) {
    for (int64_t i = 0; i < n; ++i) y[i] = orc_logf(x[i]);
}

ORC_API void orc_activate(int64_t n, const float *log_scale, const float *raw_quat,
                          const float *opacity_logit, float *scale, float *quat,
                          float *opacity) {
    /* exp / normalize / sigmoid: strategies/base_gaussian_model.py:42-50 */
    for (int64_t i = 0; i < n; ++i) {
        if (log_scale)
            for (int k = 0; k < 3; ++k) scale[3 * i + k] = orc_expf(log_scale[3 * i + k]);
        if (raw_quat) {
            const float *q = raw_quat + 4 * i;
            float nn = sqrtf(((q[0] * q[0] + q[1] * q[1]) + q[2] * q[2]) + q[3] * q[3]);
            float d = fmaxf(nn, 1e-12f); /* F.normalize eps */
            for (int k = 0; k < 4; ++k) quat[4 * i + k] = q[k] / d;
        }
        if (opacity_logit) opacity[i] = orc_sigmoidf(opacity_logit[i]);
    }
}

/* ------------------------------------------------------------------------ */
/* Projection math.  CSRC/utils.cuh:10-73,146-172,234-312                    */
/* Matrices here are row-major m[r][c].                                      */
/* ------------------------------------------------------------------------ */
static void quat_to_rotmat(const float *quat, float R[3][3]) {
    /* CSRC/utils.cuh:11-29 (wxyz, normalised inside) */
    float w = quat[0], x = quat[1], y = quat[2], z = quat[3];
    float inv_norm = 1.0f / sqrtf(((x * x + y * y) + z * z) + w * w);
    x *= inv_norm; y *= inv_norm; z *= inv_norm; w *= inv_norm;
    float x2 = x * x, y2 = y * y, z2 = z * z;
    float xy = x * y, xz = x * z, yz = y * z;
    float wx = w * x, wy = w * y, wz = w * z;
    R[0][0] = 1.f - 2.f * (y2 + z2); R[0][1] = 2.f * (xy - wz); R[0][2] = 2.f * (xz + wy);
    R[1][0] = 2.f * (xy + wz); R[1][1] = 1.f - 2.f * (x2 + z2); R[1][2] = 2.f * (yz - wx);
    R[2][0] = 2.f * (xz - wy); R[2][1] = 2.f * (yz + wx); R[2][2] = 1.f - 2.f * (x2 + y2);
}

static void quat_scale_to_covar(const float *quat, const float *scale, float covar[3][3]) {
    /* CSRC/utils.cuh:55-65: M = R*S, covar = M*M^T */
    float R[3][3], M[3][3];
    quat_to_rotmat(quat, R);
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) M[r][c] = R[r][c] * scale[c];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c)
            covar[r][c] = (M[r][0] * M[c][0] + M[r][1] * M[c][1]) + M[r][2] * M[c][2];
}

typedef struct {
    float mean_c[3];
    float mean2d[2];
    float cov2d[2][2]; /* after blur */
    float det;         /* after blur */
    float conic[3];
    float compensation;
    float radius;
} proj_t;

/* returns 1 when the Gaussian survives every test; radius_rule: 0 = dense
 * (CSRC/fully_fused_projection_fwd.cu:100-102), 1 = packed
 * (CSRC/fully_fused_projection_packed_fwd.cu:117-119). */
static int project_one(const float *mean, const float *quat, const float *scale,
                       const float *viewmat, const float *K, int width, int height,
                       float eps2d, float near_plane, float far_plane, float radius_clip,
                       int radius_rule, proj_t *o) {
    float R[3][3], t[3];
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) R[r][c] = viewmat[r * 4 + c];
        t[r] = viewmat[r * 4 + 3];
    }
    /* pos_world_to_cam: CSRC/utils.cuh:235-239 */
    for (int r = 0; r < 3; ++r)
        o->mean_c[r] = ((R[r][0] * mean[0] + R[r][1] * mean[1]) + R[r][2] * mean[2]) + t[r];
    if (o->mean_c[2] < near_plane || o->mean_c[2] > far_plane) return 0;

    float covar[3][3], T1[3][3], cc[3][3];
    quat_scale_to_covar(quat, scale, covar);
    /* covar_world_to_cam: CSRC/utils.cuh:257-261  covar_c = (R*covar)*R^T */
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            T1[i][j] = (R[i][0] * covar[0][j] + R[i][1] * covar[1][j]) + R[i][2] * covar[2][j];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            cc[i][j] = (T1[i][0] * R[j][0] + T1[i][1] * R[j][1]) + T1[i][2] * R[j][2];

    /* persp_proj: CSRC/utils.cuh:146-172 */
    float fx = K[0], fy = K[4], cx = K[2], cy = K[5];
    float x = o->mean_c[0], y = o->mean_c[1], z = o->mean_c[2];
    float tan_fovx = 0.5f * (float)width / fx;
    float tan_fovy = 0.5f * (float)height / fy;
    float lim_x = 1.3f * tan_fovx, lim_y = 1.3f * tan_fovy;
    float rz = 1.f / z, rz2 = rz * rz;
    float tx = z * fminf(lim_x, fmaxf(-lim_x, x * rz));
    float ty = z * fminf(lim_y, fmaxf(-lim_y, y * rz));
    float J00 = fx * rz, J11 = fy * rz;
    float J02 = (-fx * tx) * rz2, J12 = (-fy * ty) * rz2;
    /* (J*cov)*J^T with the structural zeros dropped (exact for finite input) */
    float T00 = J00 * cc[0][0] + J02 * cc[2][0];
    float T01 = J00 * cc[0][1] + J02 * cc[2][1];
    float T02 = J00 * cc[0][2] + J02 * cc[2][2];
    float T10 = J11 * cc[1][0] + J12 * cc[2][0];
    float T11 = J11 * cc[1][1] + J12 * cc[2][1];
    float T12 = J11 * cc[1][2] + J12 * cc[2][2];
    float c00 = T00 * J00 + T02 * J02;
    float c01 = T01 * J11 + T02 * J12; /* row 0, col 1 */
    float c10 = T10 * J00 + T12 * J02; /* row 1, col 0 */
    float c11 = T11 * J11 + T12 * J12;
    o->mean2d[0] = (fx * x) * rz + cx;
    o->mean2d[1] = (fy * y) * rz + cy;

    /* add_blur: CSRC/utils.cuh:303-312 */
    float det_orig = c00 * c11 - c10 * c01;
    c00 += eps2d; c11 += eps2d;
    float det = c00 * c11 - c10 * c01;
    o->compensation = sqrtf(fmaxf(0.f, det_orig / det));
    o->det = det;
    o->cov2d[0][0] = c00; o->cov2d[0][1] = c01; o->cov2d[1][0] = c10; o->cov2d[1][1] = c11;
    if (det <= 0.f) return 0;
    /* inverse: CSRC/utils.cuh:282-294 (conic.y comes from glm M[0][1] = row1,col0) */
    float inv_det = 1.f / det;
    o->conic[0] = c11 * inv_det;
    o->conic[1] = -c10 * inv_det;
    o->conic[2] = c00 * inv_det;

    float b = 0.5f * (c00 + c11);
    float radius;
    if (radius_rule == 1) {
        float v1 = b + sqrtf(fmaxf(0.1f, b * b - det));
        float v2 = b - sqrtf(fmaxf(0.1f, b * b - det));
        radius = ceilf(3.f * sqrtf(fmaxf(v1, v2)));
    } else {
        float v1 = b + sqrtf(fmaxf(0.01f, b * b - det));
        radius = ceilf(3.f * sqrtf(v1));
    }
    o->radius = radius;
    if (radius <= radius_clip) return 0;
    if (o->mean2d[0] + radius <= 0 || o->mean2d[0] - radius >= (float)width ||
        o->mean2d[1] + radius <= 0 || o->mean2d[1] - radius >= (float)height)
        return 0;
    return 1;
}

/* S1 frustum cull.  CSRC/fully_fused_projection_packed_fwd.cu:52-131 (keep
 * tests) and strategies/base_engine.py:18-76 (only the ids are kept).
 * mask[c*N + i] = 1 iff Gaussian i is kept for camera c.  Inputs are the
 * ACTIVATED scale / rotation of all N (ENG:370-374). */
ORC_API void orc_cull_mask(int64_t N, int B, const float *means, const float *scales,
                           const float *quats, const float *viewmats, const float *Ks,
                           int width, int height, float eps2d, float near_plane,
                           float far_plane, float radius_clip, uint8_t *mask) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < N; ++i) {
        for (int c = 0; c < B; ++c) {
            proj_t o;
            mask[(int64_t)c * N + i] = (uint8_t)project_one(
                means + 3 * i, quats + 4 * i, scales + 3 * i, viewmats + 16 * c, Ks + 9 * c,
                width, height, eps2d, near_plane, far_plane, radius_clip, 1, &o);
        }
    }
}

/* S3 dense projection.  CSRC/fully_fused_projection_fwd.cu:19-129.
 * radii==0 rows leave the other outputs untouched (reference: uninitialised). */
ORC_API void orc_projection_fwd(int C, int64_t N, const float *means, const float *quats,
                                const float *scales, const float *viewmats, const float *Ks,
                                int width, int height, float eps2d, float near_plane,
                                float far_plane, float radius_clip, int32_t *radii,
                                float *means2d, float *depths, float *conics,
                                float *compensations) {
#pragma omp parallel for schedule(static)
    for (int64_t idx = 0; idx < (int64_t)C * N; ++idx) {
        int c = (int)(idx / N);
        int64_t i = idx % N;
        proj_t o;
        int ok = project_one(means + 3 * i, quats + 4 * i, scales + 3 * i, viewmats + 16 * c,
                             Ks + 9 * c, width, height, eps2d, near_plane, far_plane,
                             radius_clip, 0, &o);
        if (!ok) { radii[idx] = 0; continue; }
        radii[idx] = (int32_t)o.radius;
        means2d[2 * idx] = o.mean2d[0];
        means2d[2 * idx + 1] = o.mean2d[1];
        depths[idx] = o.mean_c[2];
        conics[3 * idx] = o.conic[0];
        conics[3 * idx + 1] = o.conic[1];
        conics[3 * idx + 2] = o.conic[2];
        if (compensations) compensations[idx] = o.compensation;
    }
}

/* ------------------------------------------------------------------------ */
/* small row-major 3x3 helpers for the VJP (no bit-exactness needed there)   */
/* ------------------------------------------------------------------------ */
static void mm3(const float A[3][3], const float B[3][3], float C[3][3]) {
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            C[i][j] = A[i][0] * B[0][j] + A[i][1] * B[1][j] + A[i][2] * B[2][j];
}
static void tr3(const float A[3][3], float B[3][3]) {
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) B[i][j] = A[j][i];
}

/* S3 backward.  CSRC/fully_fused_projection_bwd.cu:19-185 with the VJPs of
 * CSRC/utils.cuh:31-52 (quat), :75-107 (covar), :174-232 (persp), :241-280
 * (world->cam), :296-301 (inverse).  Accumulates (+=) into v_means [N,3],
 * v_quats [N,4], v_scales [N,3] which the caller zeroes.  v_compensations is
 * not used by CLM and is not restated. */
ORC_API void orc_projection_bwd(int C, int64_t N, const float *means, const float *quats,
                                const float *scales, const float *viewmats, const float *Ks,
                                int width, int height, const int32_t *radii,
                                const float *conics, const float *v_means2d,
                                const float *v_depths, const float *v_conics, float *v_means,
                                float *v_quats, float *v_scales) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < N; ++i)
    for (int c = 0; c < C; ++c) {
        int64_t idx = (int64_t)c * N + i;
        if (radii[idx] <= 0) continue;
        const float *vm = viewmats + 16 * c, *K = Ks + 9 * c;
        const float *mean = means + 3 * i, *quat = quats + 4 * i, *scale = scales + 3 * i;
        float R[3][3], t[3];
        for (int r = 0; r < 3; ++r) {
            for (int cc = 0; cc < 3; ++cc) R[r][cc] = vm[r * 4 + cc];
            t[r] = vm[r * 4 + 3];
        }
        /* inverse_vjp: v_cov2d = -P * v_P * P, P = conic (symmetric) */
        float P[2][2] = {{conics[3 * idx], conics[3 * idx + 1]},
                         {conics[3 * idx + 1], conics[3 * idx + 2]}};
        float vP[2][2] = {{v_conics[3 * idx], v_conics[3 * idx + 1] * .5f},
                          {v_conics[3 * idx + 1] * .5f, v_conics[3 * idx + 2]}};
        float tmp[2][2], vcov2d[2][2];
        for (int a = 0; a < 2; ++a)
            for (int b = 0; b < 2; ++b) tmp[a][b] = P[a][0] * vP[0][b] + P[a][1] * vP[1][b];
        for (int a = 0; a < 2; ++a)
            for (int b = 0; b < 2; ++b)
                vcov2d[a][b] = -(tmp[a][0] * P[0][b] + tmp[a][1] * P[1][b]);

        float covar[3][3], Rt[3][3], T1[3][3], cc3[3][3], mean_c[3];
        quat_scale_to_covar(quat, scale, covar);
        for (int r = 0; r < 3; ++r)
            mean_c[r] = ((R[r][0] * mean[0] + R[r][1] * mean[1]) + R[r][2] * mean[2]) + t[r];
        tr3(R, Rt);
        mm3(R, covar, T1);
        mm3(T1, Rt, cc3);

        /* persp_proj_vjp */
        float fx = K[0], fy = K[4];
        float x = mean_c[0], y = mean_c[1], z = mean_c[2];
        float lim_x = 1.3f * (0.5f * (float)width / fx), lim_y = 1.3f * (0.5f * (float)height / fy);
        float rz = 1.f / z, rz2 = rz * rz, rz3 = rz2 * rz;
        float tx = z * fminf(lim_x, fmaxf(-lim_x, x * rz));
        float ty = z * fminf(lim_y, fmaxf(-lim_y, y * rz));
        float J[2][3] = {{fx * rz, 0.f, -fx * tx * rz2}, {0.f, fy * rz, -fy * ty * rz2}};
        float v_cc[3][3]; /* J^T * vcov2d * J */
        float JtV[3][2];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 2; ++b) JtV[a][b] = J[0][a] * vcov2d[0][b] + J[1][a] * vcov2d[1][b];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) v_cc[a][b] = JtV[a][0] * J[0][b] + JtV[a][1] * J[1][b];
        float v_mean_c[3];
        v_mean_c[0] = fx * rz * v_means2d[2 * idx];
        v_mean_c[1] = fy * rz * v_means2d[2 * idx + 1];
        v_mean_c[2] = -(fx * x * v_means2d[2 * idx] + fy * y * v_means2d[2 * idx + 1]) * rz2;
        /* v_J = vcov2d * J * cov^T + vcov2d^T * J * cov   (2x3) */
        float VJ[2][3], VtJ[2][3], v_J[2][3];
        for (int a = 0; a < 2; ++a)
            for (int b = 0; b < 3; ++b) {
                VJ[a][b] = vcov2d[a][0] * J[0][b] + vcov2d[a][1] * J[1][b];
                VtJ[a][b] = vcov2d[0][a] * J[0][b] + vcov2d[1][a] * J[1][b];
            }
        for (int a = 0; a < 2; ++a)
            for (int b = 0; b < 3; ++b)
                v_J[a][b] = (VJ[a][0] * cc3[b][0] + VJ[a][1] * cc3[b][1] + VJ[a][2] * cc3[b][2]) +
                            (VtJ[a][0] * cc3[0][b] + VtJ[a][1] * cc3[1][b] + VtJ[a][2] * cc3[2][b]);
        if (x * rz <= lim_x && x * rz >= -lim_x) v_mean_c[0] += -fx * rz2 * v_J[0][2];
        else v_mean_c[2] += -fx * rz3 * v_J[0][2] * tx;
        if (y * rz <= lim_y && y * rz >= -lim_y) v_mean_c[1] += -fy * rz2 * v_J[1][2];
        else v_mean_c[2] += -fy * rz3 * v_J[1][2] * ty;
        v_mean_c[2] += -fx * rz2 * v_J[0][0] - fy * rz2 * v_J[1][1] +
                       2.f * fx * tx * rz3 * v_J[0][2] + 2.f * fy * ty * rz3 * v_J[1][2];
        v_mean_c[2] += v_depths[idx];

        /* world->cam VJPs: v_mean = R^T v_mean_c ; v_covar = R^T v_cc R */
        float v_mean[3];
        for (int a = 0; a < 3; ++a)
            v_mean[a] = R[0][a] * v_mean_c[0] + R[1][a] * v_mean_c[1] + R[2][a] * v_mean_c[2];
        float T2[3][3], v_covar[3][3];
        mm3(Rt, v_cc, T2);
        mm3(T2, R, v_covar);

        /* quat_scale_to_covar_vjp */
        float Rq[3][3], S[3] = {scale[0], scale[1], scale[2]}, M[3][3];
        quat_to_rotmat(quat, Rq);
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) M[a][b] = Rq[a][b] * S[b];
        float sym[3][3], v_M[3][3], v_R[3][3];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) sym[a][b] = v_covar[a][b] + v_covar[b][a];
        mm3(sym, M, v_M);
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) v_R[a][b] = v_M[a][b] * S[b];
        float v_scale[3];
        for (int b = 0; b < 3; ++b)
            v_scale[b] = Rq[0][b] * v_M[0][b] + Rq[1][b] * v_M[1][b] + Rq[2][b] * v_M[2][b];
        /* quat_to_rotmat_vjp: CSRC/utils.cuh:31-52 (v_R there is glm col-major:
         * glm v_R[c][r] == v_R[r][c] here) */
        float w = quat[0], qx = quat[1], qy = quat[2], qz = quat[3];
        float inv_norm = 1.0f / sqrtf(qx * qx + qy * qy + qz * qz + w * w);
        qx *= inv_norm; qy *= inv_norm; qz *= inv_norm; w *= inv_norm;
#define VR(c, r) v_R[r][c]
        float vq[4];
        vq[0] = 2.f * (qx * (VR(1, 2) - VR(2, 1)) + qy * (VR(2, 0) - VR(0, 2)) + qz * (VR(0, 1) - VR(1, 0)));
        vq[1] = 2.f * (-2.f * qx * (VR(1, 1) + VR(2, 2)) + qy * (VR(0, 1) + VR(1, 0)) +
                       qz * (VR(0, 2) + VR(2, 0)) + w * (VR(1, 2) - VR(2, 1)));
        vq[2] = 2.f * (qx * (VR(0, 1) + VR(1, 0)) - 2.f * qy * (VR(0, 0) + VR(2, 2)) +
                       qz * (VR(1, 2) + VR(2, 1)) + w * (VR(2, 0) - VR(0, 2)));
        vq[3] = 2.f * (qx * (VR(0, 2) + VR(2, 0)) + qy * (VR(1, 2) + VR(2, 1)) -
                       2.f * qz * (VR(0, 0) + VR(1, 1)) + w * (VR(0, 1) - VR(1, 0)));
#undef VR
        float qn[4] = {w, qx, qy, qz};
        float dotq = vq[0] * qn[0] + vq[1] * qn[1] + vq[2] * qn[2] + vq[3] * qn[3];
        for (int k = 0; k < 4; ++k) v_quats[4 * i + k] += (vq[k] - dotq * qn[k]) * inv_norm;
        for (int k = 0; k < 3; ++k) {
            v_means[3 * i + k] += v_mean[k];
            v_scales[3 * i + k] += v_scale[k];
        }
    }
}

/* ------------------------------------------------------------------------ */
/* Spherical harmonics.  CSRC/spherical_harmonics.cuh:11-91 (forward) and    */
/* CK/compute_sh_bwd.cu:14-273 (backward, accumulating += into v_coeffs).    */
/* Restated as basis values b[k] and their partials db[k]/d{x,y,z}.          */
/* ------------------------------------------------------------------------ */
static void sh_basis(int degree, float x, float y, float z, float *b, float *dx, float *dy,
                     float *dz) {
    int nb = (degree + 1) * (degree + 1);
    if (dx) for (int k = 0; k < nb; ++k) dx[k] = dy[k] = dz[k] = 0.f;
    b[0] = 0.2820947917738781f;
    if (degree < 1) return;
    b[1] = -0.48860251190292f * y; b[2] = 0.48860251190292f * z; b[3] = -0.48860251190292f * x;
    if (dx) { dy[1] = -0.48860251190292f; dz[2] = 0.48860251190292f; dx[3] = -0.48860251190292f; }
    if (degree < 2) return;
    float z2 = z * z;
    float fTmp0B = -1.092548430592079f * z;
    float fC1 = x * x - y * y, fS1 = 2.f * x * y;
    b[4] = 0.5462742152960395f * fS1; b[5] = fTmp0B * y;
    b[6] = 0.9461746957575601f * z2 - 0.3153915652525201f;
    b[7] = fTmp0B * x; b[8] = 0.5462742152960395f * fC1;
    float fC1_x = 2.f * x, fC1_y = -2.f * y, fS1_x = 2.f * y, fS1_y = 2.f * x;
    if (dx) {
        dx[4] = 0.5462742152960395f * fS1_x; dy[4] = 0.5462742152960395f * fS1_y;
        dy[5] = fTmp0B; dz[5] = -1.092548430592079f * y;
        dz[6] = 2.f * 0.9461746957575601f * z;
        dx[7] = fTmp0B; dz[7] = -1.092548430592079f * x;
        dx[8] = 0.5462742152960395f * fC1_x; dy[8] = 0.5462742152960395f * fC1_y;
    }
    if (degree < 3) return;
    float fTmp0C = -2.285228997322329f * z2 + 0.4570457994644658f;
    float fTmp1B = 1.445305721320277f * z;
    float fC2 = x * fC1 - y * fS1, fS2 = x * fS1 + y * fC1;
    b[9] = -0.5900435899266435f * fS2; b[10] = fTmp1B * fS1; b[11] = fTmp0C * y;
    b[12] = z * (1.865881662950577f * z2 - 1.119528997770346f);
    b[13] = fTmp0C * x; b[14] = fTmp1B * fC1; b[15] = -0.5900435899266435f * fC2;
    float fC2_x = fC1 + x * fC1_x - y * fS1_x, fC2_y = x * fC1_y - fS1 - y * fS1_y;
    float fS2_x = fS1 + x * fS1_x + y * fC1_x, fS2_y = x * fS1_y + fC1 + y * fC1_y;
    float fTmp0C_z = -2.285228997322329f * 2.f * z;
    float b12_z = 3.f * 1.865881662950577f * z2 - 1.119528997770346f;
    if (dx) {
        dx[9] = -0.5900435899266435f * fS2_x; dy[9] = -0.5900435899266435f * fS2_y;
        dx[10] = fTmp1B * fS1_x; dy[10] = fTmp1B * fS1_y; dz[10] = 1.445305721320277f * fS1;
        dy[11] = fTmp0C; dz[11] = fTmp0C_z * y;
        dz[12] = b12_z;
        dx[13] = fTmp0C; dz[13] = fTmp0C_z * x;
        dx[14] = fTmp1B * fC1_x; dy[14] = fTmp1B * fC1_y; dz[14] = 1.445305721320277f * fC1;
        dx[15] = -0.5900435899266435f * fC2_x; dy[15] = -0.5900435899266435f * fC2_y;
    }
    if (degree < 4) return;
    float fTmp0D = z * (-4.683325804901025f * z2 + 2.007139630671868f);
    float fTmp1C = 3.31161143515146f * z2 - 0.47308734787878f;
    float fTmp2B = -1.770130769779931f * z;
    float fC3 = x * fC2 - y * fS2, fS3 = x * fS2 + y * fC2;
    b[16] = 0.6258357354491763f * fS3; b[17] = fTmp2B * fS2; b[18] = fTmp1C * fS1;
    b[19] = fTmp0D * y;
    b[20] = 1.984313483298443f * z * b[12] - 1.006230589874905f * b[6];
    b[21] = fTmp0D * x; b[22] = fTmp1C * fC1; b[23] = fTmp2B * fC2;
    b[24] = 0.6258357354491763f * fC3;
    if (dx) {
        float fTmp0D_z = 3.f * -4.683325804901025f * z2 + 2.007139630671868f;
        float fTmp1C_z = 2.f * 3.31161143515146f * z;
        float fC3_x = fC2 + x * fC2_x - y * fS2_x, fC3_y = x * fC2_y - fS2 - y * fS2_y;
        float fS3_x = fS2 + y * fC2_x + x * fS2_x, fS3_y = x * fS2_y + fC2 + y * fC2_y;
        dx[16] = 0.6258357354491763f * fS3_x; dy[16] = 0.6258357354491763f * fS3_y;
        dx[17] = fTmp2B * fS2_x; dy[17] = fTmp2B * fS2_y; dz[17] = -1.770130769779931f * fS2;
        dx[18] = fTmp1C * fS1_x; dy[18] = fTmp1C * fS1_y; dz[18] = fTmp1C_z * fS1;
        dy[19] = fTmp0D; dz[19] = fTmp0D_z * y;
        dz[20] = 1.984313483298443f * (b[12] + z * b12_z) - 1.006230589874905f * dz[6];
        dx[21] = fTmp0D; dz[21] = fTmp0D_z * x;
        dx[22] = fTmp1C * fC1_x; dy[22] = fTmp1C * fC1_y; dz[22] = fTmp1C_z * fC1;
        dx[23] = fTmp2B * fC2_x; dy[23] = fTmp2B * fC2_y; dz[23] = -1.770130769779931f * fC2;
        dx[24] = 0.6258357354491763f * fC3_x; dy[24] = 0.6258357354491763f * fC3_y;
    }
}

/* colors[n,3] = sum_k b_k(dir/|dir|) * coeffs[n,k,3]   (CSRC/compute_sh_fwd.cu:10-29) */
ORC_API void orc_sh_fwd(int64_t N, int K, int degree, const float *dirs, const float *coeffs,
                        const uint8_t *masks, float *colors) {
    int nb = (degree + 1) * (degree + 1);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < N; ++i) {
        if (masks && !masks[i]) continue;
        const float *d = dirs + 3 * i;
        float x = 0, y = 0, z = 0, b[25];
        if (degree >= 1) {
            float inorm = 1.0f / sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
            x = d[0] * inorm; y = d[1] * inorm; z = d[2] * inorm;
        }
        sh_basis(degree, x, y, z, b, NULL, NULL, NULL);
        for (int c = 0; c < 3; ++c) {
            float r = 0.f;
            for (int k = 0; k < nb; ++k) r += b[k] * coeffs[(i * K + k) * 3 + c];
            colors[3 * i + c] = r;
        }
    }
}

/* v_coeffs[n,k,3] += b_k * v_colors ; v_dirs[n,3] = projected gradient.
 * CK/compute_sh_bwd.cu:276-306 (+=, caller-owned v_coeffs) */
ORC_API void orc_sh_bwd(int64_t N, int K, int degree, const float *dirs, const float *coeffs,
                        const uint8_t *masks, const float *v_colors, float *v_coeffs,
                        float *v_dirs) {
    int nb = (degree + 1) * (degree + 1);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < N; ++i) {
        if (v_dirs) v_dirs[3 * i] = v_dirs[3 * i + 1] = v_dirs[3 * i + 2] = 0.f;
        if (masks && !masks[i]) continue;
        const float *d = dirs + 3 * i;
        float x = 0, y = 0, z = 0, inorm = 0, b[25], dx[25], dy[25], dz[25];
        if (degree >= 1) {
            inorm = 1.0f / sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
            x = d[0] * inorm; y = d[1] * inorm; z = d[2] * inorm;
        }
        sh_basis(degree, x, y, z, b, dx, dy, dz);
        float vx = 0, vy = 0, vz = 0;
        for (int c = 0; c < 3; ++c) {
            float vc = v_colors[3 * i + c];
            for (int k = 0; k < nb; ++k) {
                v_coeffs[(i * K + k) * 3 + c] += b[k] * vc;
                float cv = coeffs[(i * K + k) * 3 + c] * vc;
                vx += dx[k] * cv; vy += dy[k] * cv; vz += dz[k] * cv;
            }
        }
        if (v_dirs && degree >= 1) {
            float dot = vx * x + vy * y + vz * z;
            v_dirs[3 * i] = (vx - dot * x) * inorm;
            v_dirs[3 * i + 1] = (vy - dot * y) * inorm;
            v_dirs[3 * i + 2] = (vz - dot * z) * inorm;
        }
    }
}

/* ------------------------------------------------------------------------ */
/* S4 tile intersection.  CSRC/isect_tiles.cu:15-98 (kernel), :100-232 (host) */
/* ------------------------------------------------------------------------ */
static int32_t clampi(float v, int32_t hi) {
    /* (uint32_t)floor(negative) saturates to 0 on the GPU; clamp in the signed
     * domain like the torch twin (_torch_impl.py:223-228). */
    if (!(v > 0.f)) return 0;
    if (v >= (float)hi) return hi;
    return (int32_t)v;
}

static void tile_rect(const float *mean2d, int32_t radius_i, int tile_size, int tw, int th,
                      int32_t *x0, int32_t *y0, int32_t *x1, int32_t *y1) {
    float radius = (float)radius_i;
    float tile_radius = radius / (float)tile_size;
    float tile_x = mean2d[0] / (float)tile_size;
    float tile_y = mean2d[1] / (float)tile_size;
    *x0 = clampi(floorf(tile_x - tile_radius), tw);
    *y0 = clampi(floorf(tile_y - tile_radius), th);
    *x1 = clampi(ceilf(tile_x + tile_radius), tw);
    *y1 = clampi(ceilf(tile_y + tile_radius), th);
}

ORC_API int orc_tile_n_bits(int n_tiles) {
    /* (uint32_t)floor(log2(n_tiles)) + 1, CSRC/isect_tiles.cu:143 */
    int b = 0;
    while ((1 << b) <= n_tiles) ++b;
    return b;
}

/* pass 1: tiles_per_gauss; returns n_isects */
ORC_API int64_t orc_isect_count(int C, int64_t N, const float *means2d, const int32_t *radii,
                                int tile_size, int tw, int th, int32_t *tiles_per_gauss) {
    int64_t total = 0;
    for (int64_t idx = 0; idx < (int64_t)C * N; ++idx) {
        if (radii[idx] <= 0) { tiles_per_gauss[idx] = 0; continue; }
        int32_t x0, y0, x1, y1;
        tile_rect(means2d + 2 * idx, radii[idx], tile_size, tw, th, &x0, &y0, &x1, &y1);
        tiles_per_gauss[idx] = (y1 - y0) * (x1 - x0);
        total += tiles_per_gauss[idx];
    }
    return total;
}

static void radix_sort_pairs(int64_t n, uint64_t *keys, int32_t *vals, int end_bit) {
    /* stable LSD radix sort, 8 bits per pass = semantics of
     * cub::DeviceRadixSort::SortPairs(begin_bit=0, end_bit) CSRC/isect_tiles.cu:196-203 */
    uint64_t *k2 = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(n ? n : 1));
    int32_t *v2 = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n ? n : 1));
    uint64_t *ks = keys, *kd = k2;
    int32_t *vs = vals, *vd = v2;
    for (int shift = 0; shift < end_bit; shift += 8) {
        int bits = end_bit - shift < 8 ? end_bit - shift : 8;
        uint64_t maskb = (1ull << bits) - 1;
        int64_t hist[257];
        memset(hist, 0, sizeof(hist));
        for (int64_t i = 0; i < n; ++i) hist[((ks[i] >> shift) & maskb) + 1]++;
        for (int b = 0; b < 256; ++b) hist[b + 1] += hist[b];
        for (int64_t i = 0; i < n; ++i) {
            int64_t p = hist[(ks[i] >> shift) & maskb]++;
            kd[p] = ks[i]; vd[p] = vs[i];
        }
        uint64_t *tk = ks; ks = kd; kd = tk;
        int32_t *tv = vs; vs = vd; vd = tv;
    }
    if (ks != keys) { memcpy(keys, ks, sizeof(uint64_t) * (size_t)n); memcpy(vals, vs, sizeof(int32_t) * (size_t)n); }
    free(k2); free(v2);
}

/* pass 2 + sort: isect_ids [n_isects] int64, flatten_ids [n_isects] int32 */
ORC_API void orc_isect_emit_sort(int C, int64_t N, const float *means2d, const int32_t *radii,
                                 const float *depths, int tile_size, int tw, int th, int sort,
                                 int64_t *isect_ids, int32_t *flatten_ids) {
    int tile_n_bits = orc_tile_n_bits(tw * th);
    int cam_n_bits = orc_tile_n_bits(C); /* floor(log2(C))+1 */
    int64_t cur = 0;
    for (int64_t idx = 0; idx < (int64_t)C * N; ++idx) {
        if (radii[idx] <= 0) continue;
        int32_t x0, y0, x1, y1;
        tile_rect(means2d + 2 * idx, radii[idx], tile_size, tw, th, &x0, &y0, &x1, &y1);
        int64_t cid = idx / N;
        int64_t cid_enc = cid << (32 + tile_n_bits);
        int32_t dbits;
        memcpy(&dbits, depths + idx, 4);
        int64_t depth_enc = (int64_t)dbits; /* sign-extends like the reference cast */
        for (int32_t i = y0; i < y1; ++i)
            for (int32_t j = x0; j < x1; ++j) {
                int64_t tile_id = (int64_t)i * tw + j;
                isect_ids[cur] = cid_enc | (tile_id << 32) | depth_enc;
                flatten_ids[cur] = (int32_t)idx;
                ++cur;
            }
    }
    if (sort) radix_sort_pairs(cur, (uint64_t *)isect_ids, flatten_ids, 32 + tile_n_bits + cam_n_bits);
}

/* CSRC/isect_tiles.cu:234-278: first index of each (camera, tile) */
ORC_API void orc_isect_offset_encode(int64_t n_isects, const int64_t *isect_ids, int C,
                                     int tw, int th, int32_t *offsets) {
    int n_tiles = tw * th;
    int tile_n_bits = orc_tile_n_bits(n_tiles);
    int64_t total = (int64_t)C * n_tiles;
    if (n_isects == 0) { for (int64_t i = 0; i < total; ++i) offsets[i] = 0; return; }
    int64_t next = 0; /* next offsets slot to fill */
    for (int64_t idx = 0; idx < n_isects; ++idx) {
        int64_t hi = isect_ids[idx] >> 32;
        int64_t id = (hi >> tile_n_bits) * n_tiles + (hi & ((1 << tile_n_bits) - 1));
        while (next <= id) offsets[next++] = (int32_t)idx;
    }
    while (next < total) offsets[next++] = (int32_t)n_isects;
}

/* ------------------------------------------------------------------------ */
/* S5 rasterise.  CSRC/rasterize_to_pixels_fwd.cu:88-159                     */
/* ------------------------------------------------------------------------ */
ORC_API void orc_rasterize_fwd(int C, int64_t N, int64_t n_isects, const float *means2d,
                               const float *conics, const float *colors,
                               const float *opacities, const float *backgrounds, int W, int H,
                               int tile_size, int tw, int th, const int32_t *tile_offsets,
                               const int32_t *flatten_ids, float *render_colors,
                               float *render_alphas, int32_t *last_ids, int64_t *examined) {
    (void)N;
    int64_t exam_total = 0;
#pragma omp parallel for schedule(dynamic, 8) reduction(+ : exam_total)
    for (int64_t pix = 0; pix < (int64_t)C * H * W; ++pix) {
        int cam = (int)(pix / ((int64_t)H * W));
        int64_t p = pix % ((int64_t)H * W);
        int i = (int)(p / W), j = (int)(p % W);
        int tile_id = (i / tile_size) * tw + (j / tile_size);
        int64_t tglob = (int64_t)cam * tw * th + tile_id;
        int32_t start = tile_offsets[tglob];
        int32_t end = (tglob == (int64_t)C * tw * th - 1) ? (int32_t)n_isects : tile_offsets[tglob + 1];
        float px = (float)j + 0.5f, py = (float)i + 0.5f;
        float T = 1.0f, out[3] = {0, 0, 0};
        int32_t cur_idx = 0;
        for (int32_t idx = start; idx < end; ++idx) {
            ++exam_total;
            int32_t g = flatten_ids[idx];
            float dx = means2d[2 * g] - px, dy = means2d[2 * g + 1] - py;
            float ca = conics[3 * g], cb = conics[3 * g + 1], cc = conics[3 * g + 2];
            float sigma = 0.5f * (ca * dx * dx + cc * dy * dy) + cb * dx * dy;
            float alpha = fminf(0.999f, opacities[g] * expf(-sigma));
            if (sigma < 0.f || alpha < 1.f / 255.f) continue;
            float next_T = T * (1.0f - alpha);
            if (next_T <= 1e-4f) break; /* exclusive termination */
            float vis = alpha * T;
            for (int k = 0; k < 3; ++k) out[k] += colors[3 * g + k] * vis;
            cur_idx = idx;
            T = next_T;
        }
        render_alphas[pix] = 1.0f - T;
        for (int k = 0; k < 3; ++k)
            render_colors[3 * pix + k] = backgrounds ? out[k] + T * backgrounds[3 * cam + k] : out[k];
        last_ids[pix] = cur_idx;
    }
    if (examined) *examined = exam_total;
}

/* CSRC/rasterize_to_pixels_bwd.cu:105-248.  Gradients are accumulated in
 * double here (the oracle is the more accurate side; the GPU sums fp32 in a
 * different order, covered by the 1e-3 relative tolerance).  Outputs are
 * overwritten. */
ORC_API void orc_rasterize_bwd(int C, int64_t N, int64_t n_isects, const float *means2d,
                               const float *conics, const float *colors,
                               const float *opacities, const float *backgrounds, int W, int H,
                               int tile_size, int tw, int th, const int32_t *tile_offsets,
                               const int32_t *flatten_ids, const float *render_alphas,
                               const int32_t *last_ids, const float *v_render_colors,
                               const float *v_render_alphas, float *v_means2d_abs,
                               float *v_means2d, float *v_conics, float *v_colors,
                               float *v_opacities) {
    int64_t CN = (int64_t)C * N;
    double *acc = (double *)calloc((size_t)CN * 11, sizeof(double));
#define ACC(p, v) do { _Pragma("omp atomic") (p) += (v); } while (0)
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t pix = 0; pix < (int64_t)C * H * W; ++pix) {
        int cam = (int)(pix / ((int64_t)H * W));
        int64_t p = pix % ((int64_t)H * W);
        int i = (int)(p / W), j = (int)(p % W);
        int tile_id = (i / tile_size) * tw + (j / tile_size);
        int64_t tglob = (int64_t)cam * tw * th + tile_id;
        int32_t start = tile_offsets[tglob];
        float px = (float)j + 0.5f, py = (float)i + 0.5f;
        float T_final = 1.0f - render_alphas[pix];
        float T = T_final;
        float buffer[3] = {0, 0, 0};
        int32_t bin_final = last_ids[pix];
        const float *vrc = v_render_colors + 3 * pix;
        float vra = v_render_alphas ? v_render_alphas[pix] : 0.f;
        int32_t end = (tglob == (int64_t)C * tw * th - 1) ? (int32_t)n_isects : tile_offsets[tglob + 1];
        int32_t top = bin_final < end - 1 ? bin_final : end - 1;
        for (int32_t idx = top; idx >= start; --idx) {
            int32_t g = flatten_ids[idx];
            float dx = means2d[2 * g] - px, dy = means2d[2 * g + 1] - py;
            float ca = conics[3 * g], cb = conics[3 * g + 1], cc = conics[3 * g + 2];
            float opac = opacities[g];
            float sigma = 0.5f * (ca * dx * dx + cc * dy * dy) + cb * dx * dy;
            float vis = expf(-sigma);
            float alpha = fminf(0.999f, opac * vis);
            if (sigma < 0.f || alpha < 1.f / 255.f) continue;
            float ra = 1.0f / (1.0f - alpha);
            T *= ra;
            float fac = alpha * T;
            double *a = acc + (int64_t)g * 11;
            float v_alpha = 0.f;
            for (int k = 0; k < 3; ++k) {
                ACC(a[5 + k], (double)(fac * vrc[k]));
                v_alpha += (colors[3 * g + k] * T - buffer[k] * ra) * vrc[k];
            }
            v_alpha += T_final * ra * vra;
            if (backgrounds) {
                float accum = 0.f;
                for (int k = 0; k < 3; ++k) accum += backgrounds[3 * cam + k] * vrc[k];
                v_alpha += -T_final * ra * accum;
            }
            if (opac * vis <= 0.999f) {
                float v_sigma = -opac * vis * v_alpha;
                ACC(a[2], (double)(0.5f * v_sigma * dx * dx));
                ACC(a[3], (double)(v_sigma * dx * dy));
                ACC(a[4], (double)(0.5f * v_sigma * dy * dy));
                float gx = v_sigma * (ca * dx + cb * dy), gy = v_sigma * (cb * dx + cc * dy);
                ACC(a[0], (double)gx); ACC(a[1], (double)gy);
                ACC(a[9], (double)fabsf(gx)); ACC(a[10], (double)fabsf(gy));
                ACC(a[8], (double)(vis * v_alpha));
            }
            for (int k = 0; k < 3; ++k) buffer[k] += colors[3 * g + k] * fac;
        }
    }
    for (int64_t g = 0; g < CN; ++g) {
        const double *a = acc + g * 11;
        v_means2d[2 * g] = (float)a[0]; v_means2d[2 * g + 1] = (float)a[1];
        v_conics[3 * g] = (float)a[2]; v_conics[3 * g + 1] = (float)a[3]; v_conics[3 * g + 2] = (float)a[4];
        v_colors[3 * g] = (float)a[5]; v_colors[3 * g + 1] = (float)a[6]; v_colors[3 * g + 2] = (float)a[7];
        v_opacities[g] = (float)a[8];
        if (v_means2d_abs) { v_means2d_abs[2 * g] = (float)a[9]; v_means2d_abs[2 * g + 1] = (float)a[10]; }
    }
    free(acc);
}

/* ------------------------------------------------------------------------ */
/* Loss.  strategies/base_engine.py:79-94 + CK/ssim.cu:187-366               */
/* image: [3,H,W] float (CHW, as the engine permutes it, ENG:119);           */
/* gt_u8: [3,H,W] uint8.  Returns loss; v_image (dL/dimage, CHW) optional.   */
/* ------------------------------------------------------------------------ */
static const float GK[11] = {0.001028380123898387f, 0.0075987582094967365f, 0.036000773310661316f,
                             0.10936068743467331f, 0.21300552785396576f, 0.26601171493530273f,
                             0.21300552785396576f, 0.10936068743467331f, 0.036000773310661316f,
                             0.0075987582094967365f, 0.001028380123898387f};

static void conv11(const float *in, float *out, float *tmp, int H, int W) {
    /* zero "same" padding, x then y (CK/ssim.cu:104-185) */
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            float v = 0.f;
            for (int k = 0; k < 11; ++k) {
                int xx = x + k - 5;
                if (xx >= 0 && xx < W) v += GK[k] * in[(int64_t)y * W + xx];
            }
            tmp[(int64_t)y * W + x] = v;
        }
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            float v = 0.f;
            for (int k = 0; k < 11; ++k) {
                int yy = y + k - 5;
                if (yy >= 0 && yy < H) v += GK[k] * tmp[(int64_t)yy * W + x];
            }
            out[(int64_t)y * W + x] = v;
        }
}

ORC_API float orc_loss_fwd_bwd(int H, int W, const float *image, const uint8_t *gt_u8,
                               float lambda_dssim, float *v_image, float *ssim_mean_out,
                               float *l1_mean_out) {
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    int64_t HW = (int64_t)H * W, n = 3 * HW;
    float *buf = (float *)malloc(sizeof(float) * (size_t)HW * 12);
    float *gt = buf, *t0 = buf + HW, *tmp = buf + 2 * HW, *mu1 = buf + 3 * HW, *mu2 = buf + 4 * HW,
          *s11 = buf + 5 * HW, *s22 = buf + 6 * HW, *s12 = buf + 7 * HW, *dmu = buf + 8 * HW,
          *ds1 = buf + 9 * HW, *ds12 = buf + 10 * HW, *cv = buf + 11 * HW;
    double ssim_sum = 0.0, l1_sum = 0.0;
    float dmap = -lambda_dssim / (float)n; /* d loss / d ssim_map element */
    for (int ch = 0; ch < 3; ++ch) {
        const float *im = image + ch * HW;
        for (int64_t p = 0; p < HW; ++p) {
            float g = (float)gt_u8[ch * HW + p] / 255.0f;
            gt[p] = fminf(fmaxf(g, 0.f), 1.f);
            l1_sum += fabs((double)im[p] - (double)gt[p]);
        }
        conv11(im, mu1, tmp, H, W);
        conv11(gt, mu2, tmp, H, W);
        for (int64_t p = 0; p < HW; ++p) t0[p] = im[p] * im[p];
        conv11(t0, s11, tmp, H, W);
        for (int64_t p = 0; p < HW; ++p) t0[p] = gt[p] * gt[p];
        conv11(t0, s22, tmp, H, W);
        for (int64_t p = 0; p < HW; ++p) t0[p] = im[p] * gt[p];
        conv11(t0, s12, tmp, H, W);
        for (int64_t p = 0; p < HW; ++p) {
            float m1 = mu1[p], m2 = mu2[p];
            float sig1 = s11[p] - m1 * m1, sig2 = s22[p] - m2 * m2, sig12 = s12[p] - m1 * m2;
            float Cc = 2.0f * m1 * m2 + C1, D = 2.0f * sig12 + C2;
            float A = m1 * m1 + m2 * m2 + C1, B = sig1 + sig2 + C2;
            ssim_sum += (double)((Cc * D) / (A * B));
            dmu[p] = dmap * ((m2 * 2.0f * D) / (A * B) - (m2 * 2.0f * Cc) / (A * B) -
                             (m1 * 2.0f * Cc * D) / (A * A * B) + (m1 * 2.0f * Cc * D) / (A * B * B));
            ds1[p] = dmap * ((-Cc * D) / (A * B * B));
            ds12[p] = dmap * ((2.0f * Cc) / (A * B));
        }
        if (v_image) {
            float *vi = v_image + ch * HW;
            float l1w = (1.0f - lambda_dssim) / (float)n;
            conv11(dmu, cv, tmp, H, W);
            for (int64_t p = 0; p < HW; ++p) vi[p] = cv[p];
            conv11(ds1, cv, tmp, H, W);
            for (int64_t p = 0; p < HW; ++p) vi[p] += im[p] * 2.0f * cv[p];
            conv11(ds12, cv, tmp, H, W);
            for (int64_t p = 0; p < HW; ++p) {
                vi[p] += gt[p] * cv[p];
                float d = im[p] - gt[p];
                vi[p] += l1w * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));
            }
        }
    }
    free(buf);
    float ssim_mean = (float)(ssim_sum / (double)n), l1 = (float)(l1_sum / (double)n);
    if (ssim_mean_out) *ssim_mean_out = ssim_mean;
    if (l1_mean_out) *l1_mean_out = l1;
    return (1.0f - lambda_dssim) * l1 + lambda_dssim * (1.0f - ssim_mean);
}

/* ------------------------------------------------------------------------ */
/* Optimisers                                                                */
/* ------------------------------------------------------------------------ */
/* CK/adam.cu:12-41: masked Adam, NO bias correction */
ORC_API void orc_selective_adam(float *param, const float *grad, float *exp_avg,
                                float *exp_avg_sq, const uint8_t *mask, float lr, float b1,
                                float b2, float eps, int64_t N, int M) {
    for (int64_t p = 0; p < N * M; ++p) {
        if (!mask[p / M]) continue;
        float g = grad[p];
        float m = b1 * exp_avg[p] + (1.0f - b1) * g;
        float v = b2 * exp_avg_sq[p] + (1.0f - b2) * g * g;
        param[p] += -lr * m / (sqrtf(v) + eps);
        exp_avg[p] = m; exp_avg_sq[p] = v;
    }
}

/* torch.optim.Adam(fused=True) semantics with bias correction (optimizer.py:127-129).
 * step is the 1-based step count AFTER increment. */
ORC_API void orc_adam_dense(float *param, const float *grad, float *exp_avg, float *exp_avg_sq,
                            float lr, float b1, float b2, float eps, int64_t step, int64_t n) {
    double bc1 = 1.0 - pow((double)b1, (double)step), bc2 = 1.0 - pow((double)b2, (double)step);
    float step_size = (float)((double)lr / bc1), bc2_sqrt = (float)sqrt(bc2);
    for (int64_t p = 0; p < n; ++p) {
        float g = grad[p];
        float m = exp_avg[p] + (g - exp_avg[p]) * (1.0f - b1);
        float v = b2 * exp_avg_sq[p] + (1.0f - b2) * g * g;
        float denom = sqrtf(v) / bc2_sqrt + eps;
        param[p] -= step_size * (m / denom);
        exp_avg[p] = m; exp_avg_sq[p] = v;
    }
}

/* cpu-adam (source missing: TarzanZhao/cpu-adam @ f169c36; parity unpinned).
 * Semantics from call sites optimizer.py:130-142, ENG:320-327, GM:161-211:
 * DeepSpeed-style Adam, bias_correction on, adamw off, wd 0, per-column lr,
 * grad pre-multiplied by `scale`, grads zeroed in place afterwards (version 3).
 * rows: the Gaussian ids to update; cols = 48; col_lr[48]. */
ORC_API void orc_cpu_adam_rows(float *param, float *grad, float *exp_avg, float *exp_avg_sq,
                               const int32_t *rows, int64_t n_rows, int cols,
                               const float *col_lr, float b1, float b2, float eps, int64_t step,
                               float scale, int zero_grad) {
    double bc1 = 1.0 - pow((double)b1, (double)step), bc2 = 1.0 - pow((double)b2, (double)step);
    float inv_bc1 = (float)(1.0 / bc1), bc2_rsqrt = (float)(1.0 / sqrt(bc2));
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < n_rows; ++r) {
        int64_t base = (int64_t)rows[r] * cols;
        for (int c = 0; c < cols; ++c) {
            float g = grad[base + c] * scale;
            float m = b1 * exp_avg[base + c] + (1.0f - b1) * g;
            float v = b2 * exp_avg_sq[base + c] + (1.0f - b2) * g * g;
            float denom = sqrtf(v) * bc2_rsqrt + eps;
            param[base + c] += -(col_lr[c] * inv_bc1) * (m / denom);
            exp_avg[base + c] = m; exp_avg_sq[base + c] = v;
            if (zero_grad) grad[base + c] = 0.f;
        }
    }
}
