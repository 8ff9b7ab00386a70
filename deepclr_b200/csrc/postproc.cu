// Post-processing either side of the forward, batched on the device: pose label -> 4x4 matrix
// (LabelType.to_matrix, deepclr/data/labels.py:80-101 with _dqnormalize :46-51) and the KITTI rotation / translation
// errors between two sets of matrices (deepclr/evaluation/metrics.py:8-49).  The reference does both per pair on the
// host in float64 numpy (scripts/inference.py:113-119); here one thread per pair, float64 as well, so a batch of poses
// never has to leave the device to be compared.
#include "common.cuh"
#include "kernels.cuh"

namespace dclr {

__device__ inline void quat2mat(double w, double x, double y, double z, double *R /*3x3 row-major*/) {
    // transforms3d.quaternions.quat2mat
    const double n = w * w + x * x + y * y + z * z;
    if (n < 2.220446049250313e-16) {
        for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0) ? 1.0 : 0.0;
        return;
    }
    const double s = 2.0 / n;
    R[0] = 1 - s * (y * y + z * z); R[1] = s * (x * y - w * z);     R[2] = s * (x * z + w * y);
    R[3] = s * (x * y + w * z);     R[4] = 1 - s * (x * x + z * z); R[5] = s * (y * z - w * x);
    R[6] = s * (x * z - w * y);     R[7] = s * (y * z + w * x);     R[8] = 1 - s * (x * x + y * y);
}

__global__ void pose_to_matrix_kernel(const float *__restrict__ y, int B, int dim, int label_type, double inv_scale,
                                      double *__restrict__ T) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double l[8];
    for (int i = 0; i < dim; ++i) l[i] = (double)y[(size_t)b * dim + i] * inv_scale;
    double R[9], t[3];
    if (label_type == 0) {                       // x y z roll pitch yaw [deg], static xyz: R = Rz Ry Rx
        const double d2r = 0.017453292519943295;
        const double ai = l[3] * d2r, aj = l[4] * d2r, ak = l[5] * d2r;
        const double ci = cos(ai), si = sin(ai), cj = cos(aj), sj = sin(aj), ck = cos(ak), sk = sin(ak);
        R[0] = cj * ck; R[1] = si * sj * ck - ci * sk; R[2] = ci * sj * ck + si * sk;
        R[3] = cj * sk; R[4] = si * sj * sk + ci * ck; R[5] = ci * sj * sk - si * ck;
        R[6] = -sj;     R[7] = si * cj;                R[8] = ci * cj;
        t[0] = l[0]; t[1] = l[1]; t[2] = l[2];
    } else if (label_type == 1) {                // pos xyz | rot wxyz
        quat2mat(l[3], l[4], l[5], l[6], R);
        t[0] = l[0]; t[1] = l[1]; t[2] = l[2];
    } else {                                     // dual quaternion: real wxyz | dual wxyz
        const double nrm = sqrt(l[0] * l[0] + l[1] * l[1] + l[2] * l[2] + l[3] * l[3]) + 1e-8;
        double r[4], d[4];
        for (int i = 0; i < 4; ++i) { r[i] = l[i] / nrm; d[i] = l[4 + i] / nrm; }
        quat2mat(r[0], r[1], r[2], r[3], R);
        // t = 2 * (dual (x) conj(real))[1:]
        const double w2 = r[0], x2 = -r[1], y2 = -r[2], z2 = -r[3];
        t[0] = 2.0 * (d[0] * x2 + d[1] * w2 + d[2] * z2 - d[3] * y2);
        t[1] = 2.0 * (d[0] * y2 - d[1] * z2 + d[2] * w2 + d[3] * x2);
        t[2] = 2.0 * (d[0] * z2 + d[1] * y2 - d[2] * x2 + d[3] * w2);
    }
    double *o = T + (size_t)b * 16;
    for (int i = 0; i < 3; ++i) { o[i * 4] = R[i * 3]; o[i * 4 + 1] = R[i * 3 + 1]; o[i * 4 + 2] = R[i * 3 + 2]; o[i * 4 + 3] = t[i]; }
    o[12] = 0.0; o[13] = 0.0; o[14] = 0.0; o[15] = 1.0;
}

// rigid inverse is NOT assumed: general 4x4 affine inverse through the 3x3 block (what np.linalg.inv gives for
// [A t; 0 1]): inv = [A^-1, -A^-1 t; 0 1]
__device__ inline void affine_inv(const double *M, double *I) {
    const double a = M[0], b = M[1], c = M[2], d = M[4], e = M[5], f = M[6], g = M[8], h = M[9], i = M[10];
    const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    const double s = 1.0 / det;
    const double A[9] = {(e * i - f * h) * s, (c * h - b * i) * s, (b * f - c * e) * s,
                         (f * g - d * i) * s, (a * i - c * g) * s, (c * d - a * f) * s,
                         (d * h - e * g) * s, (b * g - a * h) * s, (a * e - b * d) * s};
    for (int r = 0; r < 3; ++r) {
        I[r * 4] = A[r * 3]; I[r * 4 + 1] = A[r * 3 + 1]; I[r * 4 + 2] = A[r * 3 + 2];
        I[r * 4 + 3] = -(A[r * 3] * M[3] + A[r * 3 + 1] * M[7] + A[r * 3 + 2] * M[11]);
    }
    I[12] = 0; I[13] = 0; I[14] = 0; I[15] = 1;
}
__device__ inline void diff_errors(const double *P, const double *Q, double &rot, double &trans) {
    // diff = P . inv(Q); rotation: acos(clamp((trace - 1) / 2)); translation: |diff[:3,3]|
    double Qi[16];
    affine_inv(Q, Qi);
    double tr = 0.0, tv[3];
    for (int r = 0; r < 3; ++r) {
        tr += P[r * 4] * Qi[r] + P[r * 4 + 1] * Qi[4 + r] + P[r * 4 + 2] * Qi[8 + r];
        tv[r] = P[r * 4] * Qi[3] + P[r * 4 + 1] * Qi[7] + P[r * 4 + 2] * Qi[11] + P[r * 4 + 3];
    }
    const double dd = 0.5 * (tr - 1.0);
    rot = acos(fmax(fmin(dd, 1.0), -1.0));
    trans = sqrt(tv[0] * tv[0] + tv[1] * tv[1] + tv[2] * tv[2]);
}

__global__ void pose_errors_kernel(const double *__restrict__ T1, const double *__restrict__ T2, int B,
                                   double *__restrict__ rot_rad, double *__restrict__ trans) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double P[16], Q[16];
    for (int i = 0; i < 16; ++i) { P[i] = T1[(size_t)b * 16 + i]; Q[i] = T2[(size_t)b * 16 + i]; }
    double r1, t1, r2, t2;
    diff_errors(P, Q, r1, t1);                    // m1 . inv(m2)
    diff_errors(Q, P, r2, t2);                    // m2 . inv(m1); the smaller of the two (metrics.py:16-20, 45-49)
    rot_rad[b] = r1 < r2 ? r1 : r2;
    trans[b] = t1 < t2 ? t1 : t2;
}

}  // namespace dclr

extern "C" int dclr_pose_to_matrix(const float *labels, int B, int label_type, float scale, double *T_out, void *stream) {
    using namespace dclr;
    DCLR_CHECK_ARG(labels && T_out, "dclr_pose_to_matrix: null pointer");
    DCLR_CHECK_ARG(B >= 0 && label_type >= 0 && label_type <= 2, "dclr_pose_to_matrix: bad arguments");
    if (B == 0) return DCLR_OK;
    const int dim = label_type == 0 ? 6 : (label_type == 1 ? 7 : 8);
    const double inv_scale = scale > 0.f ? 1.0 / (double)scale : 1.0;
    pose_to_matrix_kernel<<<div_up(B, 128), 128, 0, (cudaStream_t)stream>>>(labels, B, dim, label_type, inv_scale, T_out);
    DCLR_LAUNCH_CHECK("pose_to_matrix_kernel");
    return DCLR_OK;
}

extern "C" int dclr_pose_errors(const double *T1, const double *T2, int B, double *rot_rad_out, double *trans_out,
                                void *stream) {
    using namespace dclr;
    DCLR_CHECK_ARG(T1 && T2 && rot_rad_out && trans_out, "dclr_pose_errors: null pointer");
    DCLR_CHECK_ARG(B >= 0, "dclr_pose_errors: bad batch");
    if (B == 0) return DCLR_OK;
    pose_errors_kernel<<<div_up(B, 128), 128, 0, (cudaStream_t)stream>>>(T1, T2, B, rot_rad_out, trans_out);
    DCLR_LAUNCH_CHECK("pose_errors_kernel");
    return DCLR_OK;
}
