// C++ caller of the pipelined front end through the plain C ABI (include/gforb.h): host frames in, host results out.
// usage: frontend_cli frames.bin map.bin W H NF B K sample_scale steps out.bin
//   frames.bin: n x H x W bytes (n >= B; steps walk through it B frames at a time, wrapping)
//   map.bin   : NF x 3 float world | NF int32 score | 13 double Xv | 8 float sigma2
//   out.bin   : the LAST step's outputs, concatenated (keypoints, descriptors, counts, matches12, nmatches, good_idx, good_logdet, good_n)
// Prints one JSON line: frames/s over all steps (wall clock around gforb_frontend_sync, uploads and downloads included).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>
#include "gforb.h"

static std::vector<char> slurp(const char* path) {
    FILE* f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
    fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
    std::vector<char> b(n);
    if (fread(b.data(), 1, n, f) != (size_t)n) { fprintf(stderr, "short read %s\n", path); exit(2); }
    fclose(f);
    return b;
}
template <class T> static T* pinned(size_t n) { void* p = nullptr; if (cudaHostAlloc(&p, n * sizeof(T), cudaHostAllocDefault) != cudaSuccess) { fprintf(stderr, "cudaHostAlloc failed\n"); exit(3); } return (T*)p; }

int main(int argc, char** argv) {
    if (argc < 11) { fprintf(stderr, "usage: %s frames.bin map.bin W H NF B K sample_scale steps out.bin\n", argv[0]); return 2; }
    const int W = atoi(argv[3]), H = atoi(argv[4]), NF = atoi(argv[5]), B = atoi(argv[6]), K = atoi(argv[7]);
    const double scale = atof(argv[8]);
    const int steps = atoi(argv[9]);
    std::vector<char> frames = slurp(argv[1]), map = slurp(argv[2]);
    const size_t px = (size_t)W * H, nFrames = frames.size() / px;
    if ((int)nFrames < B) { fprintf(stderr, "need at least B frames\n"); return 2; }
    const float* world = (const float*)map.data();
    const int32_t* score = (const int32_t*)(map.data() + (size_t)NF * 12);
    const double* Xv = (const double*)(map.data() + (size_t)NF * 16);
    const float* sigma2 = (const float*)(map.data() + (size_t)NF * 16 + 13 * 8);
    GforbFrontEndConfig c;
    memset(&c, 0, sizeof(c));
    c.width = W; c.height = H; c.nfeatures = NF; c.nlevels = 8; c.scaleFactor = 1.2f; c.iniThFAST = 20; c.minThFAST = 7; c.max_batch = B;
    c.window_size = 100; c.nnratio = 0.9f; c.check_orientation = 1; c.n_good = K; c.sample_scale = scale;
    c.dx = c.dy = 0.01123325985; c.fu = 5.1369248 / c.dx; c.fv = 5.1369248 / c.dy; c.Cx = 367.215; c.Cy = 248.375;   // tools/simu_greedy.cc:33-41
    c.overlap = 1; c.device = 0;
    GforbFrontEnd* fe = nullptr;
    int rc = gforb_frontend_create(&c, &fe);
    if (rc) { fprintf(stderr, "gforb_frontend_create: %d\n", rc); return 1; }
    if ((rc = gforb_frontend_set_map(fe, world, score, Xv, sigma2))) { fprintf(stderr, "set_map: %d\n", rc); return 1; }
    // two sets of pinned staging buffers: step t+1 is enqueued while step t is still in flight
    uint8_t* himg[2]; GforbFrontEndOutputs o[2];
    for (int i = 0; i < 2; i++) {
        himg[i] = pinned<uint8_t>((size_t)B * px);
        o[i].keypoints = pinned<GforbKeyPoint>((size_t)B * NF); o[i].descriptors = pinned<uint8_t>((size_t)B * NF * 32);
        o[i].counts = pinned<int32_t>(B); o[i].matches12 = pinned<int32_t>((size_t)B * NF); o[i].nmatches = pinned<int32_t>(B);
        o[i].good_idx = pinned<int32_t>((size_t)B * K); o[i].good_logdet = pinned<double>((size_t)B * K); o[i].good_n = pinned<int32_t>(B);
    }
    auto fill = [&](int buf, int step) {
        for (int b = 0; b < B; b++) memcpy(himg[buf] + (size_t)b * px, frames.data() + (((size_t)step * B + b) % nFrames) * px, px);
    };
    fill(0, 0);
    if ((rc = gforb_frontend_step_host(fe, himg[0], B, &o[0])) || (rc = gforb_frontend_sync(fe))) { fprintf(stderr, "warm-up: %d\n", rc); return 1; }
    // timed: the frames of a step are already in pinned memory when it is issued (the copy into himg models the camera driver)
    double best = 0;
    const auto t0 = std::chrono::steady_clock::now();
    for (int s = 1; s <= steps; s++) {
        if (s >= 3 && (rc = gforb_frontend_wait_results(fe, 1))) { fprintf(stderr, "wait: %d\n", rc); return 1; }   // step s-2 has left staging set s & 1
        fill(s & 1, s);
        if ((rc = gforb_frontend_step_host(fe, himg[s & 1], B, &o[s & 1]))) { fprintf(stderr, "step %d: %d\n", s, rc); return 1; }
    }
    if ((rc = gforb_frontend_sync(fe))) { fprintf(stderr, "sync: %d\n", rc); return 1; }
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    best = (double)steps * B / sec;
    const GforbFrontEndOutputs& L = o[steps & 1];
    FILE* f = fopen(argv[10], "wb");
    if (!f) { fprintf(stderr, "cannot write %s\n", argv[10]); return 2; }
    fwrite(L.keypoints, sizeof(GforbKeyPoint), (size_t)B * NF, f); fwrite(L.descriptors, 1, (size_t)B * NF * 32, f);
    fwrite(L.counts, 4, B, f); fwrite(L.matches12, 4, (size_t)B * NF, f); fwrite(L.nmatches, 4, B, f);
    fwrite(L.good_idx, 4, (size_t)B * K, f); fwrite(L.good_logdet, 8, (size_t)B * K, f); fwrite(L.good_n, 4, B, f);
    fclose(f);
    long kp = 0, nm = 0;
    for (int b = 0; b < B; b++) { kp += L.counts[b]; nm += L.nmatches[b]; }
    printf("{\"frames_per_s\": %.1f, \"steps\": %d, \"batch\": %d, \"seconds\": %.4f, \"launches_per_step\": %d, \"keypoints_last_step\": %ld, \"matches_last_step\": %ld}\n",
           best, steps, B, sec, gforb_frontend_launches_per_step(fe), kp, nm);
    gforb_frontend_destroy(fe);
    return 0;
}
