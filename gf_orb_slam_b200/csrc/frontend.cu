// gforb_frontend_*: the batched tracking front end scheduled in C++ below the C ABI (include/gforb.h, "pipelined front end").
//
// Per step of B consecutive frames (reference call chain: Tracking::GrabImage, src/Tracking.cc:461-917):
//   main stream   halo copy (last frame of the previous step -> slot 0), ORBextractor::operator() for the B frames
//   match stream  vbPrevMatched / octave gather (k_fe_prep), ORBmatcher::SearchForInitialization(frame t-1, frame t)
//   select stream Observability::runMatrixBuilding(FRAME_INFO_MATRIX) + maxVolSelection_LazierGreedy -- waits for the matcher,
//                 like the reference where both work on the frame's matched map points
//   up / down     cudaMemcpyAsync of the frames in and of the results out (step_host), one stream each
// Keypoints, descriptors, matches, information blocks and subsets are double-buffered over steps and every hand-off is
// a CUDA event, so the extraction of step t+1 overlaps the matching, selection and result copy of step t.  With
// cfg.overlap = 0 everything runs on the main stream (the serial reference timing of bench.py).
// The operators are driven through their own C-ABI entry points (gforb_extract, gforb_search_for_initialization,
// gforb_obs_build_info_frame, gforb_obs_select) with GFORB_IN_DEVICE | GFORB_OUT_DEVICE | GFORB_ASYNC.
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "common.cuh"

namespace gforb {

// vbPrevMatched[i] = F1.mvKeysUn[i].pt (src/Tracking.cc:1003-1005), the octave column of the current frames for
// reWeightInfoMat, and the pool sizes of the selection
__global__ void __launch_bounds__(256) k_fe_prep(const GforbKeyPoint* __restrict__ kps, const int32_t* __restrict__ counts, int nf, int B,
                                                 float2* __restrict__ prev, int32_t* __restrict__ octave, int32_t* __restrict__ cntSel) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
    if (i >= nf) return;
    const GforbKeyPoint p = kps[(size_t)b * nf + i];             // slot b = the frame before frame b of this step
    prev[(size_t)b * nf + i] = make_float2(p.x, p.y);
    octave[(size_t)b * nf + i] = kps[(size_t)(b + 1) * nf + i].octave;
    if (i == 0) cntSel[b] = counts[b + 1];
}

}  // namespace gforb

using namespace gforb;

// buffer sets in flight: step t's extraction may start when the matcher, the selection and the download of step t - FE_NB are
// done.  With two sets the extraction of step t waited ~0.26 ms per batch for the selection of step t - 2 (extraction 3.09 ms,
// matcher + selection 3.05 ms behind it when they co-run); a third set removes the wait.
// GFORB_FE_NB (2 .. 4) overrides the default of 3 (A/B tool).
#define FE_NB 4                 /* array size; the number of sets in use is GforbFrontEnd::nb */
struct FeStage { std::string name; double ms = 0; int n = 0; };
struct FeRec { int stage; cudaEvent_t e0, e1; };

struct GforbFrontEnd {
    int nb = 3;                            // buffer sets in flight
    GforbFrontEndConfig cfg;
    int device = 0, B = 0, nf = 0, k = 0, launches = 0, lastBatch = 0;
    long long t = 0;                      // steps issued
    GforbExtractor* ex = nullptr;
    GforbMatcher* mt = nullptr;
    GforbObs* obs = nullptr;
    uint8_t* d_img[FE_NB] = {};
    GforbKeyPoint* d_kps[FE_NB] = {};      // (B + 1) x nf, slot 0 = halo
    uint8_t* d_desc[FE_NB] = {};
    int32_t* d_counts[FE_NB] = {};
    float* d_prev[FE_NB] = {};
    int32_t* d_m12[FE_NB] = {};
    int32_t* d_nmatch[FE_NB] = {};
    int32_t* d_octave[FE_NB] = {};
    int32_t* d_cntSel[FE_NB] = {};
    double* d_blocks[FE_NB] = {};
    double* d_obsScore[FE_NB] = {};
    int32_t* d_selIdx[FE_NB] = {};
    double* d_selLd[FE_NB] = {};
    int32_t* d_selN[FE_NB] = {};
    float* d_world = nullptr; int32_t* d_score = nullptr; uint8_t* d_valid = nullptr; double* d_Xv = nullptr; float* d_sigma2 = nullptr;
    uint32_t* d_seeds = nullptr;
    bool mapSet = false;
    cudaStream_t sMain = nullptr, sMatch = nullptr, sSel = nullptr, sUp = nullptr, sDown = nullptr;
    cudaEvent_t evExtract[FE_NB], evMatch[FE_NB], evSel[FE_NB], evH2D[FE_NB], evImgFree[FE_NB], evD2H[FE_NB];
    cudaEvent_t evHostDone[8];             // ring: step t's host frames consumed and host outputs delivered (gforb_frontend_wait_results)
    std::vector<void*> allocs;
    // profiling
    bool prof = false;
    std::vector<FeStage> stages;
    std::vector<FeRec> recs;
    std::vector<cudaEvent_t> evPool;
};

namespace {

template <class T>
int dalloc(GforbFrontEnd* F, T** p, size_t n) {
    GF_CUDA_OK(cudaMalloc((void**)p, n * sizeof(T)));
    GF_CUDA_OK(cudaMemset(*p, 0, n * sizeof(T)));
    F->allocs.push_back(*p);
    return GFORB_OK;
}

int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

}  // namespace

extern "C" {

int gforb_frontend_destroy(GforbFrontEnd* F) {
    if (!F) return GFORB_ERR_ARG;
    cudaSetDevice(F->device);
    cudaDeviceSynchronize();
    if (F->ex) gforb_extractor_destroy(F->ex);
    if (F->mt) gforb_matcher_destroy(F->mt);
    if (F->obs) gforb_obs_destroy(F->obs);
    for (void* p : F->allocs) cudaFree(p);
    for (int i = 0; i < FE_NB; i++) {
        cudaEventDestroy(F->evExtract[i]); cudaEventDestroy(F->evMatch[i]); cudaEventDestroy(F->evSel[i]);
        cudaEventDestroy(F->evH2D[i]); cudaEventDestroy(F->evImgFree[i]); cudaEventDestroy(F->evD2H[i]);
    }
    for (int i = 0; i < 8; i++) cudaEventDestroy(F->evHostDone[i]);
    for (auto& r : F->recs) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
    for (auto e : F->evPool) cudaEventDestroy(e);
    if (F->sMatch && F->sMatch != F->sMain) cudaStreamDestroy(F->sMatch);
    if (F->sSel && F->sSel != F->sMain) cudaStreamDestroy(F->sSel);
    if (F->sUp) cudaStreamDestroy(F->sUp);
    if (F->sDown) cudaStreamDestroy(F->sDown);
    if (F->sMain) cudaStreamDestroy(F->sMain);
    delete F;
    return GFORB_OK;
}

int gforb_frontend_create(const GforbFrontEndConfig* cfg, GforbFrontEnd** out) {
    if (!cfg || !out) return GFORB_ERR_ARG;
    *out = nullptr;
    if (cfg->width <= 0 || cfg->height <= 0 || cfg->nfeatures <= 0 || cfg->max_batch <= 0 || cfg->n_good <= 0) return GFORB_ERR_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || cfg->device >= ndev) return GFORB_ERR_CUDA;   // no CPU fallback
    GF_CUDA_OK(cudaSetDevice(cfg->device));
    GforbFrontEnd* F = new GforbFrontEnd();
    F->cfg = *cfg; F->device = cfg->device; F->B = cfg->max_batch; F->nf = cfg->nfeatures; F->k = cfg->n_good;
    F->nb = std::min(FE_NB, std::max(2, env_int("GFORB_FE_NB", 3)));
    for (int i = 0; i < FE_NB; i++) {
        cudaEventCreateWithFlags(&F->evExtract[i], cudaEventDisableTiming); cudaEventCreateWithFlags(&F->evMatch[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&F->evSel[i], cudaEventDisableTiming); cudaEventCreateWithFlags(&F->evH2D[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&F->evImgFree[i], cudaEventDisableTiming); cudaEventCreateWithFlags(&F->evD2H[i], cudaEventDisableTiming);
    }
    for (int i = 0; i < 8; i++) cudaEventCreateWithFlags(&F->evHostDone[i], cudaEventDisableTiming);
    int rc = GFORB_OK;
    auto fail = [&](int code) { gforb_frontend_destroy(F); return code; };
    // extraction on a high-priority stream: its short ALU-bound CTAs take freed SM slots before the long-lived CTAs of the
    // selection / matching streams do
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    const int mainPrio = std::max(hi, env_int("GFORB_MAIN_PRIORITY", -3));
    if (cudaStreamCreateWithPriority(&F->sMain, cudaStreamNonBlocking, mainPrio) != cudaSuccess) return fail(GFORB_ERR_CUDA);
    if (cfg->overlap) {
        if (cudaStreamCreateWithPriority(&F->sMatch, cudaStreamNonBlocking, std::max(hi, env_int("GFORB_MATCH_PRIORITY", 0))) != cudaSuccess) return fail(GFORB_ERR_CUDA);
        if (cudaStreamCreateWithPriority(&F->sSel, cudaStreamNonBlocking, std::max(hi, env_int("GFORB_SEL_PRIORITY", 0))) != cudaSuccess) return fail(GFORB_ERR_CUDA);
    } else {
        F->sMatch = F->sSel = F->sMain;
    }
    if (cudaStreamCreateWithFlags(&F->sUp, cudaStreamNonBlocking) != cudaSuccess) return fail(GFORB_ERR_CUDA);
    if (cudaStreamCreateWithFlags(&F->sDown, cudaStreamNonBlocking) != cudaSuccess) return fail(GFORB_ERR_CUDA);

    GforbExtractorConfig ec;
    memset(&ec, 0, sizeof(ec));
    ec.nfeatures = cfg->nfeatures; ec.scaleFactor = cfg->scaleFactor; ec.nlevels = cfg->nlevels; ec.scoreType = 1;
    ec.iniThFAST = cfg->iniThFAST; ec.minThFAST = cfg->minThFAST; ec.max_width = cfg->width; ec.max_height = cfg->height;
    ec.max_batch = cfg->max_batch; ec.device = cfg->device;
    if ((rc = gforb_extractor_create(&ec, &F->ex))) return fail(rc);
    if (!cfg->overlap) gforb_extractor_set_concurrency(F->ex, 0);      // serial reference timing: one kernel after the other
    GforbMatcherConfig mc;
    memset(&mc, 0, sizeof(mc));
    mc.max_batch = cfg->max_batch; mc.max_keypoints = cfg->nfeatures; mc.max_queries = cfg->nfeatures; mc.cand_pool = 48 * cfg->nfeatures;
    mc.nlevels = cfg->nlevels; mc.scaleFactor = cfg->scaleFactor; mc.device = cfg->device;
    if ((rc = gforb_matcher_create(&mc, &F->mt))) return fail(rc);
    GforbObsConfig oc;
    memset(&oc, 0, sizeof(oc));
    oc.fu = cfg->fu; oc.fv = cfg->fv; oc.nRows = cfg->height; oc.nCols = cfg->width; oc.Cx = cfg->Cx; oc.Cy = cfg->Cy; oc.k1 = cfg->k1; oc.k2 = cfg->k2;
    oc.dx = cfg->dx; oc.dy = cfg->dy; oc.max_batch = cfg->max_batch; oc.max_points = cfg->nfeatures; oc.device = cfg->device;
    if ((rc = gforb_obs_create(&oc, &F->obs))) return fail(rc);
    // Register budget of the selection kernel while it shares the SMs with the next step's extraction: 0 = by waves (two CTAs per
    // SM at 168 registers unless a third saves a wave), 3 = always the 128-register build.  Round 2 ran 3 until the extraction got
    // faster; alternating A/B on the last build (250-frame batches): 85.3 k -> 87.5 k frames/s device-resident with 0, end to end
    // equal (85.8 k), the matcher stage 2.7 -> 1.5 ms because its one-warp-per-frame resolver finds free slots sooner.
    if (cfg->overlap) gforb_obs_set_select_residency(F->obs, env_int("GFORB_SEL_RESIDENCY", 0));

    const size_t B = F->B, nf = F->nf, k = F->k, px = (size_t)cfg->width * cfg->height;
    for (int i = 0; i < F->nb && !rc; i++) {
        rc = dalloc(F, &F->d_img[i], B * px);
        if (!rc) rc = dalloc(F, &F->d_kps[i], (B + 1) * nf);
        if (!rc) rc = dalloc(F, &F->d_desc[i], (B + 1) * nf * 32);
        if (!rc) rc = dalloc(F, &F->d_counts[i], B + 1);
        if (!rc) rc = dalloc(F, &F->d_prev[i], B * nf * 2);
        if (!rc) rc = dalloc(F, &F->d_m12[i], B * nf);
        if (!rc) rc = dalloc(F, &F->d_nmatch[i], B);
        if (!rc) rc = dalloc(F, &F->d_octave[i], B * nf);
        if (!rc) rc = dalloc(F, &F->d_cntSel[i], B);
        if (!rc) rc = dalloc(F, &F->d_blocks[i], B * nf * 49);
        if (!rc) rc = dalloc(F, &F->d_obsScore[i], B * nf);
        if (!rc) rc = dalloc(F, &F->d_selIdx[i], B * k);
        if (!rc) rc = dalloc(F, &F->d_selLd[i], B * k);
        if (!rc) rc = dalloc(F, &F->d_selN[i], B);
    }
    if (!rc) rc = dalloc(F, &F->d_world, B * nf * 3);
    if (!rc) rc = dalloc(F, &F->d_score, B * nf);
    if (!rc) rc = dalloc(F, &F->d_valid, B * nf);
    if (!rc) rc = dalloc(F, &F->d_Xv, B * 13);
    if (!rc) rc = dalloc(F, &F->d_sigma2, (size_t)MAX_LEVELS);
    if (!rc) rc = dalloc(F, &F->d_seeds, B);
    if (rc) return fail(rc);
    std::vector<uint32_t> ones(B, 1u);                                // rand() of a process that never called srand()
    if (cudaMemcpy(F->d_seeds, ones.data(), B * 4, cudaMemcpyHostToDevice) != cudaSuccess) return fail(GFORB_ERR_CUDA);
    if (cudaMemset(F->d_valid, 1, B * nf) != cudaSuccess) return fail(GFORB_ERR_CUDA);
    *out = F;
    return GFORB_OK;
}

int gforb_frontend_set_map(GforbFrontEnd* F, const float* world, const int32_t* match_score, const double* Xv13, const float* level_sigma2) {
    if (!F || !world || !match_score || !Xv13 || !level_sigma2) return GFORB_ERR_ARG;
    GF_CUDA_OK(cudaSetDevice(F->device));
    GF_CUDA_OK(cudaDeviceSynchronize());
    const size_t nf = F->nf;
    for (int b = 0; b < F->B; b++) {
        GF_CUDA_OK(cudaMemcpy(F->d_world + (size_t)b * nf * 3, world, nf * 3 * sizeof(float), cudaMemcpyHostToDevice));
        GF_CUDA_OK(cudaMemcpy(F->d_score + (size_t)b * nf, match_score, nf * sizeof(int32_t), cudaMemcpyHostToDevice));
        GF_CUDA_OK(cudaMemcpy(F->d_Xv + (size_t)b * 13, Xv13, 13 * sizeof(double), cudaMemcpyHostToDevice));
    }
    GF_CUDA_OK(cudaMemcpy(F->d_sigma2, level_sigma2, (size_t)F->cfg.nlevels * sizeof(float), cudaMemcpyHostToDevice));
    F->mapSet = true;
    return GFORB_OK;
}

}  // extern "C"

namespace {

const int FE_FLAGS = GFORB_IN_DEVICE | GFORB_OUT_DEVICE | GFORB_ASYNC;

struct ProfScope {
    GforbFrontEnd* F; cudaStream_t st; bool on;
    ProfScope(GforbFrontEnd* F_, int stage, const char* name, cudaStream_t st_) : F(F_), st(st_), on(F_->prof && F_->recs.size() < 65536) {
        if (!on) return;
        if ((int)F->stages.size() <= stage) F->stages.resize(stage + 1);
        F->stages[stage].name = name;
        auto getEv = [&]() { cudaEvent_t e; if (!F->evPool.empty()) { e = F->evPool.back(); F->evPool.pop_back(); } else cudaEventCreate(&e); return e; };
        FeRec r; r.stage = stage; r.e0 = getEv(); r.e1 = getEv();
        F->recs.push_back(r);
        cudaEventRecord(r.e0, st);
    }
    ~ProfScope() { if (on) cudaEventRecord(F->recs.back().e1, st); }
};

// one step on frames resident at dimg; the caller has made sMain wait for whatever produces dimg
int step_common(GforbFrontEnd* F, const uint8_t* dimg, int batch) {
    const GforbFrontEndConfig& c = F->cfg;
    const int p = (int)(F->t % F->nb), nf = F->nf;
    const size_t px = (size_t)c.width * c.height;
    // buffer set p was last used by step t - FE_NB: its matcher, selection and result download must be done
    GF_CUDA_OK(cudaStreamWaitEvent(F->sMain, F->evMatch[p], 0));
    GF_CUDA_OK(cudaStreamWaitEvent(F->sMain, F->evSel[p], 0));
    GF_CUDA_OK(cudaStreamWaitEvent(F->sMain, F->evD2H[p], 0));
    int n = 0;
    {
        ProfScope ps(F, 0, "extract", F->sMain);
        if (F->t > 0) {                                               // halo: the newest frame of the previous step -> slot 0
            const int o = (p + F->nb - 1) % F->nb, s = F->lastBatch;
            GF_CUDA_OK(cudaMemcpyAsync(F->d_kps[p], F->d_kps[o] + (size_t)s * nf, (size_t)nf * sizeof(GforbKeyPoint), cudaMemcpyDeviceToDevice, F->sMain));
            GF_CUDA_OK(cudaMemcpyAsync(F->d_desc[p], F->d_desc[o] + (size_t)s * nf * 32, (size_t)nf * 32, cudaMemcpyDeviceToDevice, F->sMain));
            GF_CUDA_OK(cudaMemcpyAsync(F->d_counts[p], F->d_counts[o] + s, sizeof(int32_t), cudaMemcpyDeviceToDevice, F->sMain));
        } else {
            GF_CUDA_OK(cudaMemsetAsync(F->d_counts[p], 0, sizeof(int32_t), F->sMain));       // the first frame has no predecessor
        }
        int rc = gforb_extract(F->ex, dimg, c.width, c.height, c.width, px, batch, F->d_kps[p] + nf, F->d_desc[p] + (size_t)nf * 32,
                               F->d_counts[p] + 1, FE_FLAGS, F->sMain);
        if (rc) return rc;
        n += gforb_extractor_last_launches(F->ex);
    }
    GF_CUDA_OK(cudaEventRecord(F->evExtract[p], F->sMain));
    {
        if (F->sMatch != F->sMain) GF_CUDA_OK(cudaStreamWaitEvent(F->sMatch, F->evExtract[p], 0));
        ProfScope ps(F, 1, "match", F->sMatch);
        k_fe_prep<<<dim3((nf + 255) / 256, batch), 256, 0, F->sMatch>>>(F->d_kps[p], F->d_counts[p], nf, batch, (float2*)F->d_prev[p],
                                                                       F->d_octave[p], F->d_cntSel[p]);
        n++;
        GforbFrameSet F1 = {F->d_kps[p], F->d_desc[p], F->d_counts[p], nf, 0, c.width, 0, c.height};
        GforbFrameSet F2 = {F->d_kps[p] + nf, F->d_desc[p] + (size_t)nf * 32, F->d_counts[p] + 1, nf, 0, c.width, 0, c.height};
        int rc = gforb_search_for_initialization(F->mt, &F1, &F2, batch, F->d_prev[p], F->d_m12[p], F->d_nmatch[p], c.window_size, c.nnratio,
                                                 c.check_orientation, FE_FLAGS, F->sMatch);
        if (rc) return rc;
        n += gforb_matcher_last_launches(F->mt);
    }
    GF_CUDA_OK(cudaEventRecord(F->evMatch[p], F->sMatch));
    {
        // the good-feature stage works on the frame's matched map points: it starts after the matcher
        if (F->sSel != F->sMatch) GF_CUDA_OK(cudaStreamWaitEvent(F->sSel, F->evMatch[p], 0));
        ProfScope ps(F, 2, "info+select", F->sSel);
        int rc = gforb_obs_build_info_frame(F->obs, F->d_Xv, F->d_world, F->d_octave[p], F->d_score, F->d_valid, F->d_cntSel[p], nf, batch,
                                            F->d_sigma2, c.nlevels, F->d_blocks[p], F->d_obsScore[p], nullptr, FE_FLAGS, F->sSel);
        if (rc) return rc;
        n += gforb_obs_last_launches(F->obs);
        rc = gforb_obs_select(F->obs, F->d_blocks[p], F->d_obsScore[p], F->d_cntSel[p], nf, batch, F->k, 2, c.sample_scale, F->d_seeds,
                              F->d_selIdx[p], F->d_selLd[p], F->d_selN[p], F->k, FE_FLAGS, F->sSel);
        if (rc) return rc;
        n += gforb_obs_last_launches(F->obs);
    }
    GF_CUDA_OK(cudaEventRecord(F->evSel[p], F->sSel));
    F->launches = n;
    F->lastBatch = batch;
    F->t++;
    return GFORB_OK;
}

}  // namespace

extern "C" {

int gforb_frontend_step_device(GforbFrontEnd* F, const uint8_t* d_images, int batch) {
    if (!F || !d_images || batch <= 0 || !F->mapSet) return GFORB_ERR_ARG;
    if (batch > F->B) return GFORB_ERR_CAPACITY;
    GF_CUDA_OK(cudaSetDevice(F->device));
    return step_common(F, d_images, batch);
}

int gforb_frontend_step_host(GforbFrontEnd* F, const uint8_t* images, int batch, const GforbFrontEndOutputs* out) {
    if (!F || !images || !out || batch <= 0 || !F->mapSet) return GFORB_ERR_ARG;
    if (!out->keypoints || !out->descriptors || !out->counts || !out->matches12 || !out->nmatches || !out->good_idx || !out->good_n)
        return GFORB_ERR_ARG;
    if (batch > F->B) return GFORB_ERR_CAPACITY;
    GF_CUDA_OK(cudaSetDevice(F->device));
    const int q = (int)(F->t % F->nb), nf = F->nf, k = F->k;
    const size_t px = (size_t)F->cfg.width * F->cfg.height;
    // upload on its own stream (its own copy engine and queue): it may start as soon as the extraction that last read image
    // buffer q (step t-2) is done, i.e. while step t-1 is still computing
    GF_CUDA_OK(cudaStreamWaitEvent(F->sUp, F->evImgFree[q], 0));
    {
        ProfScope ps(F, 3, "h2d", F->sUp);
        GF_CUDA_OK(cudaMemcpyAsync(F->d_img[q], images, (size_t)batch * px, cudaMemcpyHostToDevice, F->sUp));
    }
    GF_CUDA_OK(cudaEventRecord(F->evH2D[q], F->sUp));
    GF_CUDA_OK(cudaStreamWaitEvent(F->sMain, F->evH2D[q], 0));
    int rc = step_common(F, F->d_img[q], batch);
    if (rc) return rc;
    const int p = q;                                                    // the buffer set step_common just used
    GF_CUDA_OK(cudaEventRecord(F->evImgFree[q], F->sMain));             // (sMain has only issued this step's extraction since)
    // results: keypoints as soon as the extraction is done, matches after the matcher, the subset after the selection
    GF_CUDA_OK(cudaStreamWaitEvent(F->sDown, F->evExtract[p], 0));
    {
    ProfScope psd(F, 4, "d2h keypoints+descriptors", F->sDown);
    GF_CUDA_OK(cudaMemcpyAsync(out->keypoints, F->d_kps[p] + nf, (size_t)batch * nf * sizeof(GforbKeyPoint), cudaMemcpyDeviceToHost, F->sDown));
    GF_CUDA_OK(cudaMemcpyAsync(out->descriptors, F->d_desc[p] + (size_t)nf * 32, (size_t)batch * nf * 32, cudaMemcpyDeviceToHost, F->sDown));
    GF_CUDA_OK(cudaMemcpyAsync(out->counts, F->d_counts[p] + 1, (size_t)batch * sizeof(int32_t), cudaMemcpyDeviceToHost, F->sDown));
    }
    GF_CUDA_OK(cudaStreamWaitEvent(F->sDown, F->evMatch[p], 0));
    GF_CUDA_OK(cudaMemcpyAsync(out->matches12, F->d_m12[p], (size_t)batch * nf * sizeof(int32_t), cudaMemcpyDeviceToHost, F->sDown));
    GF_CUDA_OK(cudaMemcpyAsync(out->nmatches, F->d_nmatch[p], (size_t)batch * sizeof(int32_t), cudaMemcpyDeviceToHost, F->sDown));
    GF_CUDA_OK(cudaStreamWaitEvent(F->sDown, F->evSel[p], 0));
    GF_CUDA_OK(cudaMemcpyAsync(out->good_idx, F->d_selIdx[p], (size_t)batch * k * sizeof(int32_t), cudaMemcpyDeviceToHost, F->sDown));
    if (out->good_logdet)
        GF_CUDA_OK(cudaMemcpyAsync(out->good_logdet, F->d_selLd[p], (size_t)batch * k * sizeof(double), cudaMemcpyDeviceToHost, F->sDown));
    GF_CUDA_OK(cudaMemcpyAsync(out->good_n, F->d_selN[p], (size_t)batch * sizeof(int32_t), cudaMemcpyDeviceToHost, F->sDown));
    GF_CUDA_OK(cudaEventRecord(F->evD2H[p], F->sDown));
    GF_CUDA_OK(cudaEventRecord(F->evHostDone[(F->t - 1) & 7], F->sDown));     // sDown waited for this step's upload via evExtract
    return GFORB_OK;
}

int gforb_frontend_sync(GforbFrontEnd* F) {
    if (!F) return GFORB_ERR_ARG;
    GF_CUDA_OK(cudaSetDevice(F->device));
    GF_CUDA_OK(cudaStreamSynchronize(F->sUp));
    GF_CUDA_OK(cudaStreamSynchronize(F->sMain));
    if (F->sMatch != F->sMain) GF_CUDA_OK(cudaStreamSynchronize(F->sMatch));
    if (F->sSel != F->sMain) GF_CUDA_OK(cudaStreamSynchronize(F->sSel));
    GF_CUDA_OK(cudaStreamSynchronize(F->sDown));
    int rc = gforb_extractor_poll_error(F->ex, F->sMain);
    if (!rc) rc = gforb_matcher_poll_error(F->mt, F->sMatch);
    if (!rc) rc = gforb_obs_poll_error(F->obs, F->sSel);
    return rc;
}

// host-side wait for ONE step_host call: the one issued `steps_back` (0 .. 6) calls before the most recent has consumed its host
// frames and delivered its host outputs -- what a caller cycling through S sets of staging buffers needs (steps_back = S - 1
// before refilling a set); no error poll.
int gforb_frontend_wait_results(GforbFrontEnd* F, int steps_back) {
    if (!F || steps_back < 0 || steps_back > 6 || F->t <= steps_back) return GFORB_ERR_ARG;
    GF_CUDA_OK(cudaSetDevice(F->device));
    GF_CUDA_OK(cudaEventSynchronize(F->evHostDone[(F->t - 1 - steps_back) & 7]));
    return GFORB_OK;
}

int gforb_frontend_results(GforbFrontEnd* F, GforbFrontEndOutputs* o) {
    if (!F || !o || F->t == 0) return GFORB_ERR_ARG;
    const int p = (int)((F->t - 1) % F->nb), nf = F->nf;
    o->keypoints = F->d_kps[p] + nf; o->descriptors = F->d_desc[p] + (size_t)nf * 32; o->counts = F->d_counts[p] + 1;
    o->matches12 = F->d_m12[p]; o->nmatches = F->d_nmatch[p];
    o->good_idx = F->d_selIdx[p]; o->good_logdet = F->d_selLd[p]; o->good_n = F->d_selN[p];
    return GFORB_OK;
}

int gforb_frontend_launches_per_step(const GforbFrontEnd* F) { return F ? F->launches : GFORB_ERR_ARG; }

int gforb_frontend_set_profiling(GforbFrontEnd* F, int on) {
    if (!F) return GFORB_ERR_ARG;
    F->prof = on != 0;
    for (auto& s : F->stages) { s.ms = 0; s.n = 0; }
    for (auto& r : F->recs) { F->evPool.push_back(r.e0); F->evPool.push_back(r.e1); }
    F->recs.clear();
    return gforb_extractor_set_profiling(F->ex, on);
}

// stages 0..2: extract / match / info+select (CUDA events on the stage's own stream), 3..4 after a profiled step_host: the
// frame upload and the keypoint + descriptor download; then the extractor's own stages (the first is "pyramid")
int gforb_frontend_stage_time(GforbFrontEnd* F, int i, const char** name, float* ms) {
    if (!F) return GFORB_ERR_ARG;
    GF_CUDA_OK(cudaSetDevice(F->device));
    if (!F->recs.empty()) {
        GF_CUDA_OK(cudaDeviceSynchronize());
        for (auto& r : F->recs) {
            float t = 0;
            if (cudaEventElapsedTime(&t, r.e0, r.e1) == cudaSuccess) { F->stages[r.stage].ms += t; F->stages[r.stage].n++; }
            F->evPool.push_back(r.e0); F->evPool.push_back(r.e1);
        }
        F->recs.clear();
    }
    const int own = (int)F->stages.size();
    const int nex = gforb_extractor_stage_time(F->ex, -1, nullptr, nullptr);
    const int total = own + (nex > 0 ? nex : 0);
    if (i < 0 || i >= total) return total;
    if (i < own) {
        if (name) *name = F->stages[i].name.c_str();
        if (ms) *ms = F->stages[i].n ? (float)(F->stages[i].ms / F->stages[i].n) : 0.f;
    } else {
        gforb_extractor_stage_time(F->ex, i - own, name, ms);
    }
    return total;
}

}  // extern "C"
