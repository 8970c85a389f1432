// ProblemIO.cpp -- BAL text, flat binary dump and trajectory file (see ProblemIO.h).
#include "ProblemIO.h"

#include "../../include/ba_b200.h"

#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>

namespace ba_io {
namespace {

struct FileCloser { void operator()(std::FILE *f) const { if (f) std::fclose(f); } };
typedef std::unique_ptr<std::FILE, FileCloser> File;

bool fail(std::string *err, const std::string &msg) {
    if (err) *err = msg;
    return false;
}

// Whitespace-separated numbers from a whole-file buffer: BAL files reach hundreds of MB, fscanf per token is the
// slow way.  strtod / strtol accept every spelling BAL writers emit (1e+3, -.5, 12).
struct Tokens {
    const char *p, *end;
    bool skip() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) ++p; return p < end; }
    bool real(double &v) {
        if (!skip()) return false;
        char *q; v = std::strtod(p, &q);
        if (q == p) return false;
        p = q; return true;
    }
    bool integer(long long &v) {
        if (!skip()) return false;
        char *q; v = std::strtoll(p, &q, 10);
        if (q == p) return false;
        p = q; return true;
    }
};

bool readAll(const std::string &path, std::vector<char> &buf, std::string *err) {
    File f(std::fopen(path.c_str(), "rb"));
    if (!f) return fail(err, "cannot open " + path);
    std::fseek(f.get(), 0, SEEK_END);
    const long n = std::ftell(f.get());
    std::fseek(f.get(), 0, SEEK_SET);
    if (n < 0) return fail(err, "cannot size " + path);
    buf.resize((size_t)n + 1);
    if (n > 0 && std::fread(buf.data(), 1, (size_t)n, f.get()) != (size_t)n) return fail(err, "short read on " + path);
    buf[(size_t)n] = '\0';
    return true;
}

const char MAGIC[8] = {'B', 'A', 'B', '2', '0', '0', 'P', '\1'};

template <class T>
bool putArray(std::FILE *f, const std::vector<T> &v) {
    const size_t bytes = v.size() * sizeof(T), pad = (8 - bytes % 8) % 8;
    const char zero[8] = {0};
    if (bytes && std::fwrite(v.data(), 1, bytes, f) != bytes) return false;
    return pad == 0 || std::fwrite(zero, 1, pad, f) == pad;
}
template <class T>
bool getArray(std::FILE *f, std::vector<T> &v, size_t n) {
    v.resize(n);
    const size_t bytes = n * sizeof(T), pad = (8 - bytes % 8) % 8;
    char sink[8];
    if (bytes && std::fread(v.data(), 1, bytes, f) != bytes) return false;
    return pad == 0 || std::fread(sink, 1, pad, f) == pad;
}

}  // namespace

bool FlatProblem::valid(std::string *why) const {
    const size_t Nc = cam_fixed.size(), No = obs_cam.size();
    if (cam.size() != 6 * Nc || fixed_extr.size() != 16 * Nc || K.size() != 9 * Nc) return fail(why, "camera arrays disagree in size");
    if (pt.size() % 3 != 0) return fail(why, "pt is not 3 per landmark");
    if (obs_pt.size() != No || obs_uv.size() != 2 * No) return fail(why, "observation arrays disagree in size");
    if (!bal_distortion.empty() && bal_distortion.size() != 2 * Nc) return fail(why, "bal_distortion is not 2 per camera");
    const int32_t Np = numPoints();
    for (size_t i = 0; i < No; ++i)
        if (obs_cam[i] < 0 || (size_t)obs_cam[i] >= Nc || obs_pt[i] < 0 || obs_pt[i] >= Np)
            return fail(why, "observation " + std::to_string(i) + " has an index out of range");
    return true;
}

bool readBAL(const std::string &path, FlatProblem &out, std::string *err, bool fixFirstCamera, int *n_distorted) {
    std::vector<char> buf;
    if (!readAll(path, buf, err)) return false;
    Tokens t{buf.data(), buf.data() + buf.size() - 1};
    long long Nc, Np, No;
    if (!t.integer(Nc) || !t.integer(Np) || !t.integer(No) || Nc < 0 || Np < 0 || No < 0 || Nc > INT32_MAX || Np > INT32_MAX || No > INT32_MAX)
        return fail(err, path + ": bad header (expected `cameras points observations`)");
    FlatProblem p;
    p.obs_cam.resize((size_t)No); p.obs_pt.resize((size_t)No); p.obs_uv.resize(2 * (size_t)No);
    for (long long i = 0; i < No; ++i) {
        long long c, q; double u, v;
        if (!t.integer(c) || !t.integer(q) || !t.real(u) || !t.real(v)) return fail(err, path + ": truncated at observation " + std::to_string(i));
        if (c < 0 || c >= Nc || q < 0 || q >= Np) return fail(err, path + ": observation " + std::to_string(i) + " has an index out of range");
        p.obs_cam[(size_t)i] = (int32_t)c; p.obs_pt[(size_t)i] = (int32_t)q;
        p.obs_uv[2 * (size_t)i] = (float)u; p.obs_uv[2 * (size_t)i + 1] = (float)v;      // keypoints are float32 on this path
    }
    p.cam.resize(6 * (size_t)Nc); p.cam_fixed.assign((size_t)Nc, 0); p.fixed_extr.assign(16 * (size_t)Nc, 0.f);
    p.K.assign(9 * (size_t)Nc, 0.f); p.bal_distortion.resize(2 * (size_t)Nc);
    int distorted = 0;
    for (long long c = 0; c < Nc; ++c) {
        double v[9];
        for (int k = 0; k < 9; ++k)
            if (!t.real(v[k])) return fail(err, path + ": truncated at camera " + std::to_string(c));
        for (int k = 0; k < 6; ++k) p.cam[6 * (size_t)c + k] = v[k];
        float *K = &p.K[9 * (size_t)c];
        K[0] = (float)-v[6]; K[4] = (float)-v[6]; K[8] = 1.f;          // p = -P/P.z, pixel = f p   <=>   K = diag(-f, -f, 1)
        p.bal_distortion[2 * (size_t)c] = v[7]; p.bal_distortion[2 * (size_t)c + 1] = v[8];
        if (v[7] != 0.0 || v[8] != 0.0) ++distorted;
    }
    p.pt.resize(3 * (size_t)Np);
    for (size_t k = 0; k < p.pt.size(); ++k)
        if (!t.real(p.pt[k])) return fail(err, path + ": truncated at point " + std::to_string(k / 3));
    if (fixFirstCamera && Nc > 0) {
        // the anchor's extrinsics are float32 in the reference (PointOnlyReprojectionError holds a Matrix4f, Optimizer.h:117)
        double R[9];
        ba_angle_axis_to_rotation(&p.cam[0], R);                       // column-major
        float *E = &p.fixed_extr[0];
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) E[4 * r + c] = (float)R[3 * c + r];
            E[4 * r + 3] = (float)p.cam[3 + r];
        }
        E[15] = 1.f;
        p.cam_fixed[0] = 1;
    }
    if (n_distorted) *n_distorted = distorted;
    out = std::move(p);
    return true;
}

bool writeBAL(const std::string &path, const FlatProblem &p, std::string *err) {
    std::string why;
    if (!p.valid(&why)) return fail(err, why);
    const size_t Nc = p.cam_fixed.size(), Np = (size_t)p.numPoints(), No = p.obs_cam.size();
    for (size_t c = 0; c < Nc; ++c) {
        const float *K = &p.K[9 * c];
        if (K[0] != K[4] || K[1] != 0.f || K[3] != 0.f || K[6] != 0.f || K[7] != 0.f || K[8] != 1.f || K[0] == 0.f)
            return fail(err, "camera " + std::to_string(c) + ": K is not [[f,0,cx],[0,f,cy],[0,0,1]]; BAL cannot express it");
    }
    File f(std::fopen(path.c_str(), "wb"));
    if (!f) return fail(err, "cannot open " + path + " for writing");
    std::fprintf(f.get(), "%zu %zu %zu\n", Nc, Np, No);
    for (size_t i = 0; i < No; ++i) {
        const float *K = &p.K[9 * (size_t)p.obs_cam[i]];
        // u = f Px/Pz + cx; BAL's pixel is -f_bal Px/Pz with f_bal = |f|
        const double s = K[0] < 0.f ? 1.0 : -1.0;
        const double u = s * ((double)p.obs_uv[2 * i] - (double)K[2]), v = s * ((double)p.obs_uv[2 * i + 1] - (double)K[5]);
        std::fprintf(f.get(), "%d %d %.9g %.9g\n", p.obs_cam[i], p.obs_pt[i], u, v);
    }
    for (size_t c = 0; c < Nc; ++c) {
        for (int k = 0; k < 6; ++k) std::fprintf(f.get(), "%.17g\n", p.cam[6 * c + k]);
        std::fprintf(f.get(), "%.9g\n", std::fabs((double)p.K[9 * c]));
        const double k1 = p.bal_distortion.empty() ? 0.0 : p.bal_distortion[2 * c], k2 = p.bal_distortion.empty() ? 0.0 : p.bal_distortion[2 * c + 1];
        std::fprintf(f.get(), "%.17g\n%.17g\n", k1, k2);
    }
    for (size_t k = 0; k < p.pt.size(); ++k) std::fprintf(f.get(), "%.17g\n", p.pt[k]);
    if (std::ferror(f.get())) return fail(err, "write error on " + path);
    return true;
}

bool writeBinary(const std::string &path, const FlatProblem &p, std::string *err) {
    std::string why;
    if (!p.valid(&why)) return fail(err, why);
    File f(std::fopen(path.c_str(), "wb"));
    if (!f) return fail(err, "cannot open " + path + " for writing");
    const int32_t hdr[4] = {p.numCameras(), p.numPoints(), p.numObservations(), p.bal_distortion.empty() ? 0 : 1};
    bool ok = std::fwrite(MAGIC, 1, 8, f.get()) == 8 && std::fwrite(hdr, sizeof(int32_t), 4, f.get()) == 4;
    ok = ok && putArray(f.get(), p.cam) && putArray(f.get(), p.cam_fixed) && putArray(f.get(), p.fixed_extr) && putArray(f.get(), p.K) &&
         putArray(f.get(), p.pt) && putArray(f.get(), p.obs_cam) && putArray(f.get(), p.obs_pt) && putArray(f.get(), p.obs_uv) &&
         putArray(f.get(), p.bal_distortion);
    if (!ok || std::fflush(f.get()) != 0) return fail(err, "write error on " + path);
    return true;
}

bool readBinary(const std::string &path, FlatProblem &out, std::string *err) {
    File f(std::fopen(path.c_str(), "rb"));
    if (!f) return fail(err, "cannot open " + path);
    char magic[8];
    int32_t hdr[4];
    if (std::fread(magic, 1, 8, f.get()) != 8 || std::memcmp(magic, MAGIC, 8) != 0) return fail(err, path + ": not a BAB200P file (or another version)");
    if (std::fread(hdr, sizeof(int32_t), 4, f.get()) != 4 || hdr[0] < 0 || hdr[1] < 0 || hdr[2] < 0) return fail(err, path + ": bad header");
    const size_t Nc = (size_t)hdr[0], Np = (size_t)hdr[1], No = (size_t)hdr[2];
    FlatProblem p;
    const bool ok = getArray(f.get(), p.cam, 6 * Nc) && getArray(f.get(), p.cam_fixed, Nc) && getArray(f.get(), p.fixed_extr, 16 * Nc) &&
                    getArray(f.get(), p.K, 9 * Nc) && getArray(f.get(), p.pt, 3 * Np) && getArray(f.get(), p.obs_cam, No) &&
                    getArray(f.get(), p.obs_pt, No) && getArray(f.get(), p.obs_uv, 2 * No) &&
                    getArray(f.get(), p.bal_distortion, hdr[3] ? 2 * Nc : 0);
    if (!ok) return fail(err, path + ": truncated");
    std::string why;
    if (!p.valid(&why)) return fail(err, path + ": " + why);
    out = std::move(p);
    return true;
}

int setProblem(ba_handle *h, const FlatProblem &p) {
    return ba_set_problem(h, p.numCameras(), p.cam.data(), p.cam_fixed.data(), p.fixed_extr.data(), p.K.data(), p.numPoints(),
                          p.pt.data(), p.numObservations(), p.obs_cam.data(), p.obs_pt.data(), p.obs_uv.data());
}

int fetchParameters(ba_handle *h, FlatProblem &p) {
    std::vector<double> cam(p.cam.size());
    int rc = ba_get_cameras(h, cam.data());
    if (rc != BA_OK) return rc;
    if ((rc = ba_get_points(h, p.pt.data())) != BA_OK) return rc;
    for (size_t c = 0; c < p.cam_fixed.size(); ++c)
        if (!p.cam_fixed[c]) std::memcpy(&p.cam[6 * c], &cam[6 * c], 6 * sizeof(double));
    return BA_OK;
}

void rotationToQuaternion(const float R[9], float q[4]) {
    // Eigen/src/Geometry/Quaternion.h, quaternionbase_assign_impl<Other,3,3>: "Quaternion Calculus and Fast Animation",
    // Ken Shoemake, 1987 SIGGRAPH course notes.  q = (x, y, z, w).
    auto m = [&](int r, int c) { return R[3 * r + c]; };
    float t = m(0, 0) + m(1, 1) + m(2, 2);
    if (t > 0.f) {
        t = std::sqrt(t + 1.f);
        q[3] = 0.5f * t;
        t = 0.5f / t;
        q[0] = (m(2, 1) - m(1, 2)) * t;
        q[1] = (m(0, 2) - m(2, 0)) * t;
        q[2] = (m(1, 0) - m(0, 1)) * t;
    } else {
        int i = 0;
        if (m(1, 1) > m(0, 0)) i = 1;
        if (m(2, 2) > m(i, i)) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        t = std::sqrt(m(i, i) - m(j, j) - m(k, k) + 1.f);
        q[i] = 0.5f * t;
        t = 0.5f / t;
        q[3] = (m(k, j) - m(j, k)) * t;
        q[j] = (m(j, i) + m(i, j)) * t;
        q[k] = (m(k, i) + m(i, k)) * t;
    }
}

bool writeTrajectoryTUM(const std::string &outputPath, int n, const double *timestamps, const float *poses, std::string *err) {
    const std::string path = outputPath + "_estimatedPoses.txt";
    File f(std::fopen(path.c_str(), "wb"));
    if (!f) return fail(err, "cannot open " + path + " for writing");
    for (int i = 0; i < n; ++i) {
        const float *P = poses + 16 * (size_t)i;
        const float R[9] = {P[0], P[1], P[2], P[4], P[5], P[6], P[8], P[9], P[10]};
        float q[4];
        rotationToQuaternion(R, q);
        // setprecision(4) << fixed: timestamp (double), camera centre (float), quaternion x y z w (float)
        std::fprintf(f.get(), "%.4f %.4f %.4f %.4f %.4f %.4f %.4f %.4f\n", timestamps[i], (double)P[3], (double)P[7], (double)P[11],
                     (double)q[0], (double)q[1], (double)q[2], (double)q[3]);
    }
    if (std::ferror(f.get())) return fail(err, "write error on " + path);
    return true;
}

}  // namespace ba_io
