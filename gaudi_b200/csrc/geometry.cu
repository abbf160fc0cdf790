// Geometric restraint objectives evaluated on the expressed population (SURVEY.md section 8f,
// "next" row 4): Distance (gaudi/objectives/distance.py:102-129) and Angle
// (gaudi/objectives/angle.py:86-96).  They read a handful of atoms per individual, so this is a
// trivially HBM-bound kernel: one thread per (individual, objective), each term is a distance,
// an angle or a dihedral between expressed mobile atoms and/or fixed points (static atoms or
// literal coordinates), combined the way the reference combines them; kind 3 terms accumulate a
// mass-weighted centre (Distance with center_of_mass).
#include "common.cuh"

namespace gmm {

__device__ __forceinline__ void geo_point(const GeoTerm& t, int k, const float* pose, double cx, double cy, double cz,
                                          double* p) {
    if (t.atom[k] >= 0) {          // expressed mobile atom, device frame -> world frame
        p[0] = (double)pose[3 * t.atom[k]] + cx; p[1] = (double)pose[3 * t.atom[k] + 1] + cy;
        p[2] = (double)pose[3 * t.atom[k] + 2] + cz;
    } else {                       // static atom or literal point (world frame)
        p[0] = t.point[k][0]; p[1] = t.point[k][1]; p[2] = t.point[k][2];
    }
}

__global__ void geometry_kernel(int P, int M, const float* xyz, int n_terms, const GeoTerm* terms, int n_obj,
                                double cx, double cy, double cz, double* out) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P * n_obj) return;
    const int p = idx / n_obj, obj = idx - p * n_obj;
    const float* pose = xyz + (size_t)p * M * 3;
    double sum = 0.0;
    int n = 0;
    double com[3] = {0.0, 0.0, 0.0}, mass = 0.0, com_target[3] = {0.0, 0.0, 0.0};   // kind 3
    for (int k = 0; k < n_terms; ++k) {
        const GeoTerm t = terms[k];
        if (t.obj != obj) continue;
        double a[3], b[3], c[3], d[3], v;
        geo_point(t, 0, pose, cx, cy, cz, a);
        geo_point(t, 1, pose, cx, cy, cz, b);
        if (t.kind == 3) {
            // distance.py:131-148 (center_of_mass): point 0 is a probe atom, `threshold` carries its mass,
            // point 1 the target; the objective is the distance of the mass-weighted centre to the target
            com[0] += t.threshold * a[0]; com[1] += t.threshold * a[1]; com[2] += t.threshold * a[2];
            mass += t.threshold;
            com_target[0] = b[0]; com_target[1] = b[1]; com_target[2] = b[2];
            continue;
        }
        if (t.kind == 0) {
            // distance.py:112-127: d - threshold; below `tolerance` -> -1000 * weight; mean of |.|
            const double dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
            v = sqrt(dx * dx + dy * dy + dz * dz) - t.threshold;
            if (t.has_tolerance && v < t.tolerance) v = -1000.0 * t.weight;
            v = fabs(v);
        } else {
            geo_point(t, 2, pose, cx, cy, cz, c);
            double ang;
            if (t.kind == 1) {     // chimera.angle(a, b, c): angle at b, degrees
                const double u[3] = {a[0] - b[0], a[1] - b[1], a[2] - b[2]}, w[3] = {c[0] - b[0], c[1] - b[1], c[2] - b[2]};
                const double dot = u[0] * w[0] + u[1] * w[1] + u[2] * w[2];
                const double nu = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]), nw = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
                ang = acos(fmin(1.0, fmax(-1.0, dot / (nu * nw)))) * (180.0 / 3.14159265358979323846);
            } else {               // chimera.dihedral(a, b, c, d), IUPAC sign, degrees
                geo_point(t, 3, pose, cx, cy, cz, d);
                const double b0[3] = {a[0] - b[0], a[1] - b[1], a[2] - b[2]};
                double b1[3] = {c[0] - b[0], c[1] - b[1], c[2] - b[2]};
                const double b2[3] = {d[0] - c[0], d[1] - c[1], d[2] - c[2]};
                const double n1 = sqrt(b1[0] * b1[0] + b1[1] * b1[1] + b1[2] * b1[2]);
                b1[0] /= n1; b1[1] /= n1; b1[2] /= n1;
                const double d0 = b0[0] * b1[0] + b0[1] * b1[1] + b0[2] * b1[2];
                const double d2 = b2[0] * b1[0] + b2[1] * b1[1] + b2[2] * b1[2];
                const double vv[3] = {b0[0] - d0 * b1[0], b0[1] - d0 * b1[1], b0[2] - d0 * b1[2]};
                const double ww[3] = {b2[0] - d2 * b1[0], b2[1] - d2 * b1[1], b2[2] - d2 * b1[2]};
                const double x = vv[0] * ww[0] + vv[1] * ww[1] + vv[2] * ww[2];
                const double cr[3] = {b1[1] * vv[2] - b1[2] * vv[1], b1[2] * vv[0] - b1[0] * vv[2], b1[0] * vv[1] - b1[1] * vv[0]};
                const double y = cr[0] * ww[0] + cr[1] * ww[1] + cr[2] * ww[2];
                ang = atan2(y, x) * (180.0 / 3.14159265358979323846);
            }
            // angle.py:92-95: 'planar' -> |sin(angle)|, else |threshold - angle|
            v = t.planar ? fabs(sin(ang * (3.14159265358979323846 / 180.0))) : fabs(t.threshold - ang);
        }
        sum += v;
        ++n;
    }
    if (mass > 0.0) {
        const double dx = com[0] / mass - com_target[0], dy = com[1] / mass - com_target[1], dz = com[2] / mass - com_target[2];
        out[idx] = sqrt(dx * dx + dy * dy + dz * dz);
        return;
    }
    out[idx] = n ? sum / n : 0.0;
}

int launch_geometry(gmm_ctx* c, int P, const float* xyz, double* out, cudaStream_t s) {
    if (P == 0 || c->n_geo_obj == 0) return GMM_OK;
    const int total = P * c->n_geo_obj;
    geometry_kernel<<<(total + 255) / 256, 256, 0, s>>>(P, c->M, xyz, c->n_geo_terms, c->d_geo_terms, c->n_geo_obj,
                                                      c->centre[0], c->centre[1], c->centre[2], out);
    GMM_LAUNCH_CHECK("geometry_kernel");
    return GMM_OK;
}

}  // namespace gmm
