// compat_driver.cpp — TEST: the drop-in ORB_SLAM3::ORBextractor class (orb_slam3_ros2_b200/compat) used exactly the way
// Frame::ExtractORB uses the reference (Frame.cc:418-425, accessors :110-116, pyramid reads :818), compared against
// the reference's own unmodified ORBextractor.cc (oracle/_ref/liborbref_strict.so, loaded through its C facade so the
// two identically named classes never meet in one translation unit).
//
//   compat_driver <frames.raw> <w> <h> <nframes> <nfeatures> <lap0> <lap1> <liborbref.so>
// exit 0 = every frame bit-identical (monoIndex, keypoints, descriptors, all padded pyramid planes, tables).
//   compat_driver --time <frames.raw> <w> <h> <nframes> <nfeatures> <lap0> <lap1> <reps> [<liborbref.so>]
// the DROP-IN LATENCY: median / p90 wall milliseconds of ORB_SLAM3::ORBextractor::operator() on one host thread, frame by
// frame (what Frame::ExtractORB costs the tracking thread, Frame.cc:418-425), with and without the mvImagePyramid download,
// and of the reference class on the same frames; one JSON line.
#include <dlfcn.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <vector>

#include "ORBextractor.h"

typedef void* (*ref_create_t)(int, float, int, int, int);
typedef void (*ref_destroy_t)(void*);
typedef int (*ref_extract_t)(void*, const uint8_t*, int, int, int, int, int, void*, uint8_t*, int, int*);
typedef int (*ref_pyramid_t)(void*, int, uint8_t*, int*, int*);
typedef int (*ref_tables_t)(void*, float*, float*, float*, float*, int*, int*);
typedef float (*ref_sf_t)(void*);

#define CHECK(cond, ...)                  \
    do {                                  \
        if (!(cond)) {                    \
            fprintf(stderr, __VA_ARGS__); \
            fprintf(stderr, "\n");        \
            return 1;                     \
        }                                 \
    } while (0)

static double nowMs() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}
static void stats(std::vector<double> v, double* med, double* p90) {
    std::sort(v.begin(), v.end());
    *med = v.empty() ? 0 : v[v.size() / 2];
    *p90 = v.empty() ? 0 : v[std::min(v.size() - 1, (size_t)(v.size() * 9 / 10))];
}

static int timeMode(int argc, char** argv) {
    if (argc < 10) {
        fprintf(stderr, "usage: compat_driver --time frames.raw w h nframes nfeatures lap0 lap1 reps [liborbref.so]\n");
        return 2;
    }
    const int w = atoi(argv[3]), h = atoi(argv[4]), nframes = atoi(argv[5]), nfeat = atoi(argv[6]), reps = atoi(argv[9]);
    std::vector<int> lap = {atoi(argv[7]), atoi(argv[8])};
    std::vector<uint8_t> frames((size_t)w * h * nframes);
    FILE* fp = fopen(argv[2], "rb");
    CHECK(fp, "cannot open %s", argv[2]);
    CHECK(fread(frames.data(), 1, frames.size(), fp) == frames.size(), "short read");
    fclose(fp);
    ORB_SLAM3::ORBextractor ex(nfeat, 1.2f, 8, 20, 7);
    double res[2][2] = {{0, 0}, {0, 0}};
    size_t kp = 0;
    for (int withPyr = 1; withPyr >= 0; --withPyr) {
        ex.downloadPyramid = withPyr != 0;
        std::vector<double> t;
        for (int r = -3; r < reps; ++r) {  // three warm-up calls (plan, buffers, staging)
            const int f = ((r % nframes) + nframes) % nframes;
            cv::Mat im(h, w, CV_8UC1, frames.data() + (size_t)f * w * h, (size_t)w);
            std::vector<cv::KeyPoint> keys;
            cv::Mat desc;
            const double t0 = nowMs();
            ex(im, cv::Mat(), keys, desc, lap);
            if (r >= 0) t.push_back(nowMs() - t0);
            kp = keys.size();
        }
        stats(t, &res[withPyr][0], &res[withPyr][1]);
    }
    double refMed = 0, refP90 = 0;
    int refReps = 0;
    if (argc > 10) {
        void* lib = dlopen(argv[10], RTLD_NOW | RTLD_LOCAL);
        if (lib) {
            ref_create_t rcreate = (ref_create_t)dlsym(lib, "orbref_create");
            ref_destroy_t rdestroy = (ref_destroy_t)dlsym(lib, "orbref_destroy");
            ref_extract_t rextract = (ref_extract_t)dlsym(lib, "orbref_extract");
            if (rcreate && rdestroy && rextract) {
                void* ref = rcreate(nfeat, 1.2f, 8, 20, 7);
                const int cap = nfeat + 64;
                std::vector<cv::KeyPoint> rk(cap);
                std::vector<uint8_t> rd((size_t)cap * 32);
                std::vector<double> t;
                refReps = std::min(reps, 24);
                for (int r = -1; r < refReps; ++r) {
                    const int f = ((r % nframes) + nframes) % nframes;
                    int rn = 0;
                    const double t0 = nowMs();
                    rextract(ref, frames.data() + (size_t)f * w * h, w, h, w, lap[0], lap[1], rk.data(), rd.data(), cap, &rn);
                    if (r >= 0) t.push_back(nowMs() - t0);
                }
                stats(t, &refMed, &refP90);
                rdestroy(ref);
            }
        }
    }
    printf("{\"frame\": \"%dx%d\", \"nfeatures\": %d, \"reps\": %d, \"keypoints\": %zu, \"with_pyramid_ms\": {\"median\": %.4f, \"p90\": %.4f}, "
           "\"without_pyramid_ms\": {\"median\": %.4f, \"p90\": %.4f}, \"reference_class_ms\": {\"median\": %.3f, \"p90\": %.3f, \"reps\": %d}}\n",
           w, h, nfeat, reps, kp, res[1][0], res[1][1], res[0][0], res[0][1], refMed, refP90, refReps);
    return 0;
}

int main(int argc, char** argv) {
    if (argc > 1 && strcmp(argv[1], "--time") == 0) return timeMode(argc, argv);
    if (argc < 9) {
        fprintf(stderr, "usage: compat_driver frames.raw w h nframes nfeatures lap0 lap1 liborbref.so\n");
        return 2;
    }
    const int w = atoi(argv[2]), h = atoi(argv[3]), nframes = atoi(argv[4]), nfeat = atoi(argv[5]);
    std::vector<int> lap = {atoi(argv[6]), atoi(argv[7])};
    void* lib = dlopen(argv[8], RTLD_NOW | RTLD_LOCAL);
    CHECK(lib, "dlopen %s: %s", argv[8], dlerror());
    ref_create_t rcreate = (ref_create_t)dlsym(lib, "orbref_create");
    ref_destroy_t rdestroy = (ref_destroy_t)dlsym(lib, "orbref_destroy");
    ref_extract_t rextract = (ref_extract_t)dlsym(lib, "orbref_extract");
    ref_pyramid_t rpyr = (ref_pyramid_t)dlsym(lib, "orbref_pyramid");
    ref_tables_t rtab = (ref_tables_t)dlsym(lib, "orbref_tables");
    ref_sf_t rsf = (ref_sf_t)dlsym(lib, "orbref_scale_factor");
    CHECK(rcreate && rdestroy && rextract && rpyr && rtab && rsf, "reference facade lacks a symbol");

    std::vector<uint8_t> frames((size_t)w * h * nframes);
    FILE* fp = fopen(argv[1], "rb");
    CHECK(fp, "cannot open %s", argv[1]);
    CHECK(fread(frames.data(), 1, frames.size(), fp) == frames.size(), "short read");
    fclose(fp);

    // Tracking.cc:595-601 style construction (EuRoC settings)
    ORB_SLAM3::ORBextractor* ex = new ORB_SLAM3::ORBextractor(nfeat, 1.2f, 8, 20, 7);
    void* ref = rcreate(nfeat, 1.2f, 8, 20, 7);

    // accessors (Frame.cc:110-116)
    const int nl = ex->GetLevels();
    std::vector<float> a(nl), b(nl), c(nl), d(nl);
    std::vector<int> nf(nl);
    int um[16];
    CHECK(rtab(ref, a.data(), b.data(), c.data(), d.data(), nf.data(), um) == nl, "level count differs");
    CHECK(ex->GetScaleFactor() == rsf(ref), "GetScaleFactor differs");
    CHECK(ex->GetScaleFactors() == a && ex->GetInverseScaleFactors() == b && ex->GetScaleSigmaSquares() == c &&
              ex->GetInverseScaleSigmaSquares() == d,
          "scale tables differ");

    const int cap = nfeat + 64;
    std::vector<cv::KeyPoint> rk(cap);
    std::vector<uint8_t> rd((size_t)cap * 32);
    size_t totalKp = 0;
    cv::Mat heldLevel;  // a Frame copy holding an old pyramid header must stay valid (refcount)
    std::vector<uint8_t> heldCopy;
    for (int f = 0; f < nframes; ++f) {
        cv::Mat im(h, w, CV_8UC1, frames.data() + (size_t)f * w * h, (size_t)w);
        std::vector<cv::KeyPoint> keys;
        cv::Mat desc;
        const int mono = (*ex)(im, cv::Mat(), keys, desc, lap);
        int rn = 0;
        const int rmono = rextract(ref, im.data, w, h, w, lap[0], lap[1], rk.data(), rd.data(), cap, &rn);
        CHECK(mono == rmono, "frame %d: monoIndex %d vs %d", f, mono, rmono);
        CHECK((int)keys.size() == rn, "frame %d: %zu keypoints vs %d", f, keys.size(), rn);
        CHECK(rn == 0 || memcmp(keys.data(), rk.data(), (size_t)rn * sizeof(cv::KeyPoint)) == 0, "frame %d: keypoints differ", f);
        CHECK(desc.rows == rn || (rn == 0 && desc.empty()), "frame %d: descriptor rows", f);
        for (int i = 0; i < rn; ++i)
            CHECK(memcmp(desc.ptr(i), rd.data() + (size_t)i * 32, 32) == 0, "frame %d: descriptor %d differs", f, i);
        CHECK((int)ex->mvImagePyramid.size() == nl, "pyramid size");
        for (int l = 0; l < nl; ++l) {
            int pw = 0, ph = 0;
            CHECK(rpyr(ref, l, nullptr, &pw, &ph) == 0, "ref pyramid");
            const cv::Mat& m = ex->mvImagePyramid[l];
            CHECK(m.cols == pw && m.rows == ph, "frame %d level %d: size %dx%d vs %dx%d", f, l, m.cols, m.rows, pw, ph);
            std::vector<uint8_t> plane((size_t)(pw + 38) * (ph + 38));
            rpyr(ref, l, plane.data(), nullptr, nullptr);
            const uint8_t* base = m.data - 19 * m.step - 19;  // the ROI sits at (19,19) of the padded buffer
            for (int y = 0; y < ph + 38; ++y)
                CHECK(memcmp(base + (size_t)y * m.step, plane.data() + (size_t)y * (pw + 38), pw + 38) == 0,
                      "frame %d level %d: padded row %d differs", f, l, y);
        }
        if (f == 0) {
            heldLevel = ex->mvImagePyramid[1];
            for (int y = 0; y < heldLevel.rows; ++y) heldCopy.insert(heldCopy.end(), heldLevel.ptr(y), heldLevel.ptr(y) + heldLevel.cols);
        }
        totalKp += keys.size();
    }
    // the header taken after frame 0 still shows frame 0's level 1
    for (int y = 0; y < heldLevel.rows; ++y)
        CHECK(memcmp(heldLevel.ptr(y), heldCopy.data() + (size_t)y * heldLevel.cols, heldLevel.cols) == 0,
              "held pyramid header was overwritten by a later call");

    // empty image -> -1, nothing touched (ORBextractor.cc:1090-1091)
    std::vector<cv::KeyPoint> keys(3);
    cv::Mat desc;
    CHECK((*ex)(cv::Mat(), cv::Mat(), keys, desc, lap) == -1, "empty image must return -1");
    CHECK(keys.size() == 3, "empty image must leave the keypoint vector alone");

    delete ex;
    rdestroy(ref);
    printf("compat ok: %d frames %dx%d, %zu keypoints bit-identical to the reference class\n", nframes, w, h, totalKp);
    return 0;
}
