// C ABI of the trainer level: gsb_trainer_*.  Host orchestration counterpart of Trainer::train /
// Trainer::render (reference src/Trainer.cu:148-216,252-543).
//
// Differences in mechanism (not in results):
//   * workspaces live in grow-only arenas: no cudaMalloc / cudaFree per view (T:335-337,422-424)
//   * camera constants are uploaded once per set_truths, not four blocking copies per view (T:317-326)
//   * no host synchronisation inside an iteration: R stays on the device, the binning arena is sized
//     from a high-water mark and an overflow (detected at the end of the iteration through a pinned
//     host mirror of every view's R) grows the arena and re-runs the iteration before any parameter
//     was touched
//   * loss, per-view gradient buffers, their nine memsets and accumulateGradients are fused into the
//     render-backward prologue and the preprocess-backward epilogue
//   * densify runs on the GPU (classify + scan + scatter) instead of a D2H / CPU / H2D round trip.

#include <cuda.h>
#include "workspace.cuh"
#include "trainer_kernels.cuh"
#include "../../include/gsb.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <vector>

struct gsb_trainer {
    int W = 0, H = 0, grid_x = 0, grid_y = 0, tiles = 0, tile_bits = 0;
    cudaStream_t stream = nullptr;

    // bound model (canonical storage: the caller's ModelSplatsDevice arrays)
    int count = 0, capacity = 0, D = 0, M = 0;
    float *loc = nullptr, *shs = nullptr, *scales = nullptr, *opac = nullptr, *rot = nullptr;

    // truths
    std::vector<const uint32_t*> truthW, truthB;
    std::vector<gsb_camera> cams;
    gsb_camera* d_cams = nullptr;
    size_t d_cams_cap = 0;
    gsb_camera* d_pcam = nullptr;     // camera slot of Trainer::render
    int total_cameras = 0;
    float* d_bg = nullptr;            // white rgb, black rgb

    // arenas
    char* geom = nullptr; size_t geom_P = 0; GeomWs g;
    GradRec2* grads2 = nullptr;       // scratch of a fused camera pass: (white, black) pairs
    bool fuse_wb = true;              // white + black pass of a camera share one forward / one backward walk
    char* bin = nullptr; uint32_t r_cap = 0; BinWs b;
    char* img = nullptr; ImageWs im;
    float* rendered = nullptr;        // [3][H][W]
    // The members above are the ACTIVE view workspace.  Two complete sets exist (slot[0], slot[1]) so the
    // forward chain of view pass i+1 (latency-bound, small grids) can run on a second stream underneath
    // the render backward of view pass i (issue-bound); activate() swaps a set into the members before
    // a pass is enqueued (kernel arguments are captured at launch).
    struct ViewSlot {
        char* geom = nullptr; GeomWs g; GradRec2* grads2 = nullptr;
        char* bin = nullptr; BinWs b;
        char* img = nullptr; ImageWs im;
        float* rendered = nullptr;
    } slot[2];
    int active = 0;
    bool overlap = true;              // gsb_trainer_set_overlap
    cudaStream_t sF = nullptr, sB = nullptr;          // non-blocking: forward chain / backward chain
    cudaStream_t sF2 = nullptr;                       // forward chains of the odd passes (slot 1): chains of different slots are independent
    cudaEvent_t ev_scan_prev = nullptr;               // tile scan of the second-to-last pass (it runs on the other forward stream)
    bool two_fwd_streams = true;
    bool lpt = true;                         // heavy-first tile order (GSB_LPT=0: raster order, developer A/B)
    cudaStream_t exec_f = nullptr, exec_b = nullptr;  // null = everything on `stream`
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_fdone[2] = {nullptr, nullptr}, ev_bdone[2] = {nullptr, nullptr};
    // preview arena (Trainer::render at another resolution)
    char* pimg = nullptr; size_t pimg_bytes = 0; float* pcolor = nullptr; size_t pcolor_floats = 0;
    ImageWs pim; size_t pimg_tiles = 0, pimg_N = 0;
    bool dirty = false;               // a pass was abandoned half way: tile counters / maxima are re-zeroed before the next one
    cudaEvent_t ev_x[3] = {nullptr, nullptr, nullptr};   // exchange / apply brackets (stage timing)
    cudaEvent_t ev_scan = nullptr;    // recorded after the last view pass's tile scan: every R of the iteration is on the host

    float* avg_own = nullptr; size_t avg_own_floats = 0;
    float* avg_ext = nullptr;
    uint32_t epoch = 0;
    uint32_t* h_R = nullptr; size_t h_R_cap = 0;   // pinned mirrors of every view's R

    // densify scratch
    uint8_t* flags = nullptr; uint32_t *keep_idx = nullptr, *aux_idx = nullptr, *totals = nullptr, *dens_ticket = nullptr;
    unsigned long long* dens_lb = nullptr; size_t dens_P = 0;
    bool plan_valid = false;
    gsb_project_params plan_params;
    uint32_t n_keep = 0, n_split_done = 0, n_clone_done = 0;

    gsb_allreduce_fn allreduce = nullptr; void* allreduce_user = nullptr;
    gsb_comm* comm = nullptr;         // gsb_trainer_set_comm: native exchange, takes precedence over the hook
    gsb_window* win = nullptr;        // with a communicator the running means live in a peer window (comm.cu)
    bool loss_on = false; double* d_loss = nullptr;
    // lazy read-back of the loss scalar: two pinned slots, iteration k lands in slot k & 1
    double* h_loss = nullptr; cudaEvent_t ev_loss[2] = {nullptr, nullptr}; bool loss_pending = false; unsigned loss_iter = 0;
    gsb_trainer_stats stats;

    // live stage timing (gsb_trainer_set_stage_timing): one event per stage boundary and view pass
    bool timing_on = false;
    std::vector<cudaEvent_t> ev_pool;
    double stage_ms[GSB_NUM_STAGES] = {0};
    long long stage_passes = 0;
    unsigned long long* d_pairs = nullptr;   // gsb_trainer_count_pairs scratch
};

namespace {

void save_slot(gsb_trainer* t) {
    gsb_trainer::ViewSlot& v = t->slot[t->active];
    v.geom = t->geom; v.g = t->g; v.grads2 = t->grads2; v.bin = t->bin; v.b = t->b; v.img = t->img; v.im = t->im;
    v.rendered = t->rendered;
}
void activate(gsb_trainer* t, int k) {
    if (k == t->active) return;
    save_slot(t);
    const gsb_trainer::ViewSlot& v = t->slot[k];
    t->geom = v.geom; t->g = v.g; t->grads2 = v.grads2; t->bin = v.bin; t->b = v.b; t->img = v.img; t->im = v.im;
    t->rendered = v.rendered;
    t->active = k;
}

float* avg_ptr(gsb_trainer* t) { return t->avg_ext ? t->avg_ext : t->avg_own; }
size_t grad_floats(const gsb_trainer* t) { return (size_t)t->count * (12 + 3 * (size_t)t->M); }

int ensure_geom(gsb_trainer* t, size_t P) {
    if (t->slot[0].geom && t->slot[1].geom && t->geom && P <= t->geom_P) return GSB_OK;
    GSB_CUDA_OK(cudaStreamSynchronize(t->stream));
    const size_t cap = P + P / 8 + 1024;
    const size_t bytes = GeomWs::required(cap);
    const int was = t->active;
    for (int k = 0; k < 2; k++) {
        activate(t, k);
        if (t->geom) GSB_CUDA_OK(cudaFree(t->geom));
        GSB_CUDA_OK(cudaMalloc(&t->geom, bytes));
        GSB_CUDA_OK(cudaMemsetAsync(t->geom, 0, bytes, t->stream));   // grad scratch starts at zero
        t->g.carve(t->geom, cap);
        if (t->grads2) GSB_CUDA_OK(cudaFree(t->grads2));
        GSB_CUDA_OK(cudaMalloc(&t->grads2, cap * sizeof(GradRec2)));
        GSB_CUDA_OK(cudaMemsetAsync(t->grads2, 0, cap * sizeof(GradRec2), t->stream));
        save_slot(t);
    }
    t->geom_P = cap;
    activate(t, was);
    return GSB_OK;
}

int ensure_bin(gsb_trainer* t, uint32_t r_cap) {
    if (t->bin && r_cap <= t->r_cap) return GSB_OK;
    GSB_CUDA_OK(cudaStreamSynchronize(t->stream));
    const size_t bytes = BinWs::required(r_cap);
    const int was = t->active;
    for (int k = 0; k < 2; k++) {
        activate(t, k);
        if (t->bin) GSB_CUDA_OK(cudaFree(t->bin));
        GSB_CUDA_OK(cudaMalloc(&t->bin, bytes));
        GSB_CUDA_OK(cudaMemsetAsync(t->bin, 0, bytes, t->stream));
        t->b.carve(t->bin, r_cap);
        save_slot(t);
    }
    t->r_cap = r_cap;
    activate(t, was);
    return GSB_OK;
}

int drop_window(gsb_trainer* t) {
    if (!t->win) return GSB_OK;
    const int e = gsb_comm_window_destroy(t->win);      // collective
    t->win = nullptr; t->avg_own = nullptr; t->avg_own_floats = 0;
    return e;
}

int ensure_avg(gsb_trainer* t) {
    const size_t n = grad_floats(t);
    if (t->avg_ext) return GSB_OK;
    if (t->comm && gsb_comm_world(t->comm) > 1) {
        // Collective: every rank reaches this with the same count (replicas are identical), so all of them keep or
        // replace their windows in the same iteration.
        if (t->win && n <= t->avg_own_floats) return GSB_OK;
        if (t->win) { if (int e = drop_window(t)) return e; }
        else if (t->avg_own) { GSB_CUDA_OK(cudaFree(t->avg_own)); t->avg_own = nullptr; t->avg_own_floats = 0; }
        if (int e = gsb_comm_window_create(t->comm, n + n / 2 + 256, &t->win)) return e;
        t->avg_own = gsb_window_data(t->win);
        t->avg_own_floats = gsb_window_floats(t->win);
        return GSB_OK;
    }
    if (n <= t->avg_own_floats) return GSB_OK;
    if (t->avg_own) GSB_CUDA_OK(cudaFree(t->avg_own));
    GSB_CUDA_OK(cudaMalloc(&t->avg_own, (n + n / 8 + 256) * sizeof(float)));
    t->avg_own_floats = n + n / 8 + 256;
    return GSB_OK;
}

// loss of the iteration whose partial sums (GSB_LOSS_SLOTS doubles, common.cuh) were copied into host slot `slot`
double loss_of_slot(const gsb_trainer* t, unsigned slot) {
    double v = 0.0;
    for (int k = 0; k < GSB_LOSS_SLOTS; k++) v += t->h_loss[(size_t)slot * GSB_LOSS_SLOTS + k];
    return v;
}

// Look-back status words carry a 30-bit epoch tag (binning.cu lb_pack) so the status arrays never need clearing
// between launches.  Before the tag would wrap, every stream is drained, all look-back arrays are cleared and the
// counter restarts (one view pass consumes 1 epoch, a densify 1).
constexpr uint32_t GSB_EPOCH_LIMIT = 0x3FFFFFFFu - 64u;

int epoch_guard(gsb_trainer* t) {
    if (t->epoch < GSB_EPOCH_LIMIT) return GSB_OK;
    GSB_CUDA_OK(cudaStreamSynchronize(t->stream));
    if (t->sF) GSB_CUDA_OK(cudaStreamSynchronize(t->sF));
    if (t->sF2) GSB_CUDA_OK(cudaStreamSynchronize(t->sF2));
    if (t->sB) GSB_CUDA_OK(cudaStreamSynchronize(t->sB));
    save_slot(t);
    for (int k = 0; k < 2; k++) {
        const gsb_trainer::ViewSlot& v = t->slot[k];
        if (v.img) GSB_CUDA_OK(cudaMemset(v.im.lookback, 0, ImageWs::lookback_words((size_t)t->tiles) * sizeof(unsigned long long)));
    }
    if (t->pimg) GSB_CUDA_OK(cudaMemset(t->pim.lookback, 0, ImageWs::lookback_words(t->pimg_tiles) * sizeof(unsigned long long)));
    if (t->dens_lb) GSB_CUDA_OK(cudaMemset(t->dens_lb, 0, ((t->dens_P + 1023) / 1024 + 1) * 4 * sizeof(unsigned long long)));
    t->epoch = 0;
    t->stats.epoch_wraps++;
    return GSB_OK;
}

// tile_sort_kernel leaves the per-tile counters and maxima of a workspace at zero for the next pass.  A pass that was
// abandoned half way (an error return) breaks that invariant: drain and re-zero everything before the next one.
int clean_if_dirty(gsb_trainer* t) {
    if (!t->dirty) return GSB_OK;
    GSB_CUDA_OK(cudaStreamSynchronize(t->stream));
    if (t->sF) GSB_CUDA_OK(cudaStreamSynchronize(t->sF));
    if (t->sF2) GSB_CUDA_OK(cudaStreamSynchronize(t->sF2));
    if (t->sB) GSB_CUDA_OK(cudaStreamSynchronize(t->sB));
    cudaGetLastError();
    save_slot(t);
    for (int k = 0; k < 2; k++)
        if (t->slot[k].img) GSB_CUDA_OK(cudaMemset(t->slot[k].im.tile_max, 0, t->slot[k].im.zero_bytes((size_t)t->tiles)));
    if (t->pimg) GSB_CUDA_OK(cudaMemset(t->pim.tile_max, 0, t->pim.zero_bytes(t->pimg_tiles)));
    t->dirty = false;
    return GSB_OK;
}

// Stage boundaries of one view pass (gsb_trainer_profile_pass reports the time between them).
enum { ST_CLEAR = 0, ST_PREPROCESS, ST_TILE_SCAN, ST_EMIT, ST_TILE_SORT, ST_RENDER_FWD, ST_RENDER_BWD, ST_PREPROCESS_BWD,
       ST_COUNT };
static_assert(ST_COUNT == GSB_NUM_STAGES, "stage list of include/gsb.h");

inline void mark(cudaEvent_t* ev, int k, cudaStream_t s) {
    if (ev) cudaEventRecord(ev[k], s);
}

// forward half of one view pass: preprocess (+ per-tile counts), tile scan, emit, per-tile sort, render forward
int forward_pass(gsb_trainer* t, const float* view, const float* projview, const float* campos, float tanx, float tany,
                 float scale_modifier, const float* bg, int W, int H, ImageWs& im, float* out_color, uint32_t* h_R,
                 cudaEvent_t* ev = nullptr, cudaEvent_t scan_done = nullptr, uint32_t* frame = nullptr) {
    GSB_NVTX("gsb forward chain");
    cudaStream_t s = t->exec_f ? t->exec_f : t->stream;
    const int grid_x = (W + GSB_TILE - 1) / GSB_TILE, grid_y = (H + GSB_TILE - 1) / GSB_TILE;
    const int tiles = grid_x * grid_y;
    const int P = t->count;
    if (int e = epoch_guard(t)) return e;
    mark(ev, ST_CLEAR, s);
    GSB_CUDA_OK(cudaMemsetAsync(t->g.ctrl, 0, sizeof(GsbCtrl), s));

    mark(ev, ST_PREPROCESS, s);
    PreprocessArgs pa;
    pa.P = P; pa.D = t->D; pa.M = t->M;
    pa.means3D = t->loc; pa.scales = t->scales; pa.scale_modifier = scale_modifier; pa.rotations = t->rot;
    pa.opacities = t->opac; pa.shs = t->shs; pa.cov3D_precomp = nullptr; pa.colors_precomp = nullptr;
    pa.viewmatrix = view; pa.projmatrix = projview; pa.cam_pos = campos;
    pa.W = W; pa.H = H; pa.tan_fovx = tanx; pa.tan_fovy = tany;
    pa.focal_y = H / (2.0f * tany);
    pa.focal_x = W / (2.0f * tanx);
    pa.grid_x = grid_x; pa.grid_y = grid_y; pa.prefiltered = 0;
    pa.radii = t->g.radii; pa.radii_user = nullptr; pa.depths = t->g.depths; pa.tiles_touched = t->g.tiles_touched;
    pa.recs = t->g.recs; pa.cov3D = t->g.cov3D; pa.clamped = t->g.clamped;
    pa.tile_count = im.tile_count; pa.ctrl = t->g.ctrl;
    { GSB_NVTX("preprocess"); launch_preprocess_fwd(pa, s); }

    BinningArgs ba;
    ba.P = P; ba.r_cap = t->r_cap; ba.grid_x = grid_x; ba.grid_y = grid_y;
    ba.radii = t->g.radii; ba.tiles_touched = t->g.tiles_touched; ba.depths = t->g.depths; ba.recs = t->g.recs;
    ba.tile_count = im.tile_count; ba.ranges = im.ranges; ba.tile_max = im.tile_max; ba.big_list = im.big_list;
    ba.tile_order = t->lpt ? im.tile_order : nullptr;
    ba.pairs = t->b.pairs; ba.point_list = t->b.point_list; ba.pairs2 = t->b.pairs2; ba.chunk_list = t->b.chunk_list;
    ba.lookback = im.lookback; ba.epoch = &t->epoch; ba.ctrl = t->g.ctrl; ba.num_rendered_host = h_R;
    mark(ev, ST_TILE_SCAN, s);
    { GSB_NVTX("tile scan"); launch_tile_scan(ba, s); }
    if (scan_done) GSB_CUDA_OK(cudaEventRecord(scan_done, s));   // R of this pass is in pinned host memory once this fires
    mark(ev, ST_EMIT, s);
    { GSB_NVTX("emit"); launch_emit(ba, s); }
    mark(ev, ST_TILE_SORT, s);
    { GSB_NVTX("tile sort"); launch_tile_sort(ba, s); }

    mark(ev, ST_RENDER_FWD, s);
    RenderFwdArgs ra;
    ra.W = W; ra.H = H; ra.grid_x = grid_x; ra.grid_y = grid_y;
    ra.ranges = im.ranges; ra.point_list = t->b.point_list; ra.recs = t->g.recs; ra.bg = bg;
    ra.out_color = out_color; ra.final_T = im.final_T; ra.n_contrib = im.n_contrib; ra.tile_max = im.tile_max;
    ra.frame = frame;                 // Trainer::render: RGBA8 straight from the blend registers (imageFloatToInt fused)
    ra.tile_order = ba.tile_order; ra.ctrl = t->g.ctrl;
    { GSB_NVTX("render forward"); launch_render_fwd(ra, s); }
    t->stats.kernel_launches += 6;
    GSB_CUDA_OK(cudaGetLastError());
    return GSB_OK;
}

// loss + backward + accumulate.  Unfused: training view pass i of 2n (white passes first, then black,
// T:311-314).  Fused: camera i of n, both backgrounds in one walk (the forward ran on black).
int backward_pass(gsb_trainer* t, int i, bool fused, float* avg, float S, cudaEvent_t* ev = nullptr) {
    GSB_NVTX("gsb backward chain");
    cudaStream_t s = t->exec_b ? t->exec_b : t->stream;
    const int n = (int)t->truthW.size();
    const bool white = fused || i < n;
    const int cam = i % n;
    const uint32_t* truth = white ? t->truthW[cam] : t->truthB[cam];
    const gsb_camera* dc = t->d_cams + cam;
    const float* bg = t->d_bg + ((white && !fused) ? 0 : 3);
    mark(ev, ST_RENDER_BWD, s);
    RenderBwdArgs rb;
    rb.W = t->W; rb.H = t->H; rb.grid_x = t->grid_x; rb.grid_y = t->grid_y;
    rb.ranges = t->im.ranges; rb.point_list = t->b.point_list; rb.recs = t->g.recs; rb.bg = bg;
    rb.final_T = t->im.final_T; rb.n_contrib = t->im.n_contrib; rb.tile_max = t->im.tile_max;
    rb.dL_dpix = nullptr; rb.truth = truth; rb.rendered = t->rendered; rb.grads = t->g.grads;
    rb.truth2 = fused ? t->truthB[cam] : nullptr; rb.grads2 = fused ? t->grads2 : nullptr;
    rb.loss_out = t->loss_on ? t->d_loss : nullptr;
    rb.tile_order = t->lpt ? t->im.tile_order : nullptr; rb.ctrl = t->g.ctrl;
    launch_render_bwd(rb, s);

    mark(ev, ST_PREPROCESS_BWD, s);
    PreprocessBwdArgs pb;
    memset(&pb, 0, sizeof(pb));
    pb.P = t->count; pb.D = t->D; pb.M = t->M;
    pb.means3D = t->loc; pb.scales = t->scales; pb.scale_modifier = 1.0f; pb.rotations = t->rot; pb.shs = t->shs;
    pb.cov3D = t->g.cov3D; pb.cov3D_is_precomp = 0; pb.colors_precomp = 0;
    pb.viewmatrix = dc->view; pb.projmatrix = dc->projview; pb.cam_pos = dc->location;
    pb.tan_fovx = t->cams[cam].tan_fovx; pb.tan_fovy = t->cams[cam].tan_fovy;
    pb.focal_y = t->H / (2.0f * pb.tan_fovy);
    pb.focal_x = t->W / (2.0f * pb.tan_fovx);
    pb.radii = t->g.radii; pb.clamped = t->g.clamped; pb.grads = t->g.grads; pb.recs = t->g.recs; pb.W = t->W; pb.H = t->H;
    pb.avg = avg; pb.samples = S;
    pb.grads2 = fused ? t->grads2 : nullptr;   // both passes of the camera in one sweep over the splats
    launch_preprocess_bwd(pb, s);
    t->stats.kernel_launches += 2;
    mark(ev, ST_COUNT, s);
    return GSB_OK;
}

int view_pass(gsb_trainer* t, int i, bool fused, float* avg, float S, cudaEvent_t* ev = nullptr, cudaEvent_t scan_done = nullptr) {
    const int n = (int)t->truthW.size();
    const int cam = i % n;
    const gsb_camera* dc = t->d_cams + cam;
    const float* bg = t->d_bg + ((!fused && i < n) ? 0 : 3);
    if (int e = forward_pass(t, dc->view, dc->projview, dc->location, t->cams[cam].tan_fovx, t->cams[cam].tan_fovy, 1.0f, bg,
                             t->W, t->H, t->im, t->rendered, t->h_R + i, ev, scan_done)) return e;
    if (t->exec_f) {   // pipelined: the backward chain of this pass waits for its forward chain
        GSB_CUDA_OK(cudaEventRecord(t->ev_fdone[t->active], t->exec_f));
        GSB_CUDA_OK(cudaStreamWaitEvent(t->exec_b, t->ev_fdone[t->active], 0));
    }
    if (int e = backward_pass(t, i, fused, avg, S, ev)) return e;
    if (t->exec_b) GSB_CUDA_OK(cudaEventRecord(t->ev_bdone[t->active], t->exec_b));
    return GSB_OK;
}

}  // namespace

extern "C" {

int gsb_trainer_create(gsb_trainer** out, int width, int height, void* stream) {
    if (!out || width <= 0 || height <= 0) return gsb_fail_msg("gsb_trainer_create: invalid argument");
    int ndev = 0;
    GSB_CUDA_OK(cudaGetDeviceCount(&ndev));
    if (ndev == 0) return gsb_fail_msg("gsb_trainer_create: no CUDA device (this library has no CPU path)");
    gsb_trainer* t = new gsb_trainer();
    memset(&t->stats, 0, sizeof(t->stats));
    t->W = width; t->H = height;
    t->grid_x = (width + GSB_TILE - 1) / GSB_TILE; t->grid_y = (height + GSB_TILE - 1) / GSB_TILE;
    t->tiles = t->grid_x * t->grid_y;
    t->tile_bits = gsb_tile_bits((uint32_t)t->tiles);
    t->stream = (cudaStream_t)stream;
    const float bg[6] = {1.f, 1.f, 1.f, 0.f, 0.f, 0.f};
    GSB_CUDA_OK(cudaMalloc(&t->d_bg, sizeof(bg)));
    GSB_CUDA_OK(cudaMemcpy(t->d_bg, bg, sizeof(bg), cudaMemcpyHostToDevice));
    const size_t N = (size_t)width * height;
    for (int k = 0; k < 2; k++) {
        activate(t, k);
        GSB_CUDA_OK(cudaMalloc(&t->img, ImageWs::required(N, (size_t)t->tiles)));
        t->im.carve(t->img, N, (size_t)t->tiles);
        GSB_CUDA_OK(cudaMemset(t->im.tile_max, 0, t->im.zero_bytes((size_t)t->tiles)));   // tile_sort_kernel keeps them zero between passes
        GSB_CUDA_OK(cudaMalloc(&t->rendered, 3 * N * sizeof(float)));
        save_slot(t);
    }
    activate(t, 0);
    // The forward chain is a string of small, latency-bound grids: at high priority its CTAs take whatever
    // an SM frees while the (low-priority, register-heavy) backward kernel drains.
    int prio_lo = 0, prio_hi = 0;
    GSB_CUDA_OK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    if (const char* e = getenv("GSB_STREAM_PRIO")) {      // developer A/B: 0 = equal priorities, -1 = backward high / forward low
        if (atoi(e) == 0) prio_hi = prio_lo;
        else if (atoi(e) < 0) std::swap(prio_lo, prio_hi);
    }
    GSB_CUDA_OK(cudaStreamCreateWithPriority(&t->sF, cudaStreamNonBlocking, prio_hi));
    GSB_CUDA_OK(cudaStreamCreateWithPriority(&t->sF2, cudaStreamNonBlocking, prio_hi));
    GSB_CUDA_OK(cudaStreamCreateWithPriority(&t->sB, cudaStreamNonBlocking, prio_lo));
    GSB_CUDA_OK(cudaEventCreateWithFlags(&t->ev_scan_prev, cudaEventDisableTiming));
    if (const char* e = getenv("GSB_FWD_STREAMS")) t->two_fwd_streams = atoi(e) != 1;
    if (const char* e = getenv("GSB_LPT")) t->lpt = atoi(e) != 0;
    GSB_CUDA_OK(cudaEventCreateWithFlags(&t->ev_fork, cudaEventDisableTiming));
    GSB_CUDA_OK(cudaEventCreateWithFlags(&t->ev_join, cudaEventDisableTiming));
    GSB_CUDA_OK(cudaEventCreateWithFlags(&t->ev_scan, cudaEventDisableTiming));
    for (int k = 0; k < 2; k++) {
        GSB_CUDA_OK(cudaEventCreateWithFlags(&t->ev_fdone[k], cudaEventDisableTiming));
        GSB_CUDA_OK(cudaEventCreateWithFlags(&t->ev_bdone[k], cudaEventDisableTiming));
    }
    GSB_CUDA_OK(cudaMalloc(&t->d_loss, GSB_LOSS_SLOTS * sizeof(double)));
    GSB_CUDA_OK(cudaMemset(t->d_loss, 0, GSB_LOSS_SLOTS * sizeof(double)));
    GSB_CUDA_OK(cudaHostAlloc(&t->h_loss, 2 * GSB_LOSS_SLOTS * sizeof(double), cudaHostAllocDefault));
    memset(t->h_loss, 0, 2 * GSB_LOSS_SLOTS * sizeof(double));
    for (int k = 0; k < 2; k++) GSB_CUDA_OK(cudaEventCreateWithFlags(&t->ev_loss[k], cudaEventDisableTiming));
    GSB_CUDA_OK(cudaMalloc(&t->totals, 4 * sizeof(uint32_t)));
    GSB_CUDA_OK(cudaMalloc(&t->dens_ticket, sizeof(uint32_t)));
    GSB_CUDA_OK(cudaMalloc(&t->d_pcam, sizeof(gsb_camera)));
    *out = t;
    return GSB_OK;
}

int gsb_trainer_destroy(gsb_trainer* t) {
    if (!t) return GSB_OK;
    cudaStreamSynchronize(t->stream);
    if (t->sF) { cudaStreamSynchronize(t->sF); cudaStreamDestroy(t->sF); }
    if (t->sF2) { cudaStreamSynchronize(t->sF2); cudaStreamDestroy(t->sF2); }
    if (t->ev_scan_prev) cudaEventDestroy(t->ev_scan_prev);
    if (t->sB) { cudaStreamSynchronize(t->sB); cudaStreamDestroy(t->sB); }
    save_slot(t);
    for (int k = 0; k < 2; k++) {
        cudaFree(t->slot[k].geom); cudaFree(t->slot[k].grads2); cudaFree(t->slot[k].bin); cudaFree(t->slot[k].img);
        cudaFree(t->slot[k].rendered);
        if (t->ev_fdone[k]) cudaEventDestroy(t->ev_fdone[k]);
        if (t->ev_bdone[k]) cudaEventDestroy(t->ev_bdone[k]);
    }
    if (t->ev_fork) cudaEventDestroy(t->ev_fork);
    if (t->ev_join) cudaEventDestroy(t->ev_join);
    if (t->ev_scan) cudaEventDestroy(t->ev_scan);
    for (int k = 0; k < 3; k++) if (t->ev_x[k]) cudaEventDestroy(t->ev_x[k]);
    cudaFree(t->d_cams); cudaFree(t->d_pcam); cudaFree(t->d_bg);
    if (t->win) drop_window(t);        // collective when a communicator is attached: destroy trainers before communicators
    cudaFree(t->pimg); cudaFree(t->pcolor); cudaFree(t->avg_own); cudaFree(t->d_loss);
    cudaFree(t->flags); cudaFree(t->keep_idx); cudaFree(t->aux_idx); cudaFree(t->totals); cudaFree(t->dens_ticket);
    cudaFree(t->dens_lb); cudaFree(t->d_pairs);
    for (cudaEvent_t e : t->ev_pool) cudaEventDestroy(e);
    if (t->h_R) cudaFreeHost(t->h_R);
    if (t->h_loss) cudaFreeHost(t->h_loss);
    for (int k = 0; k < 2; k++) if (t->ev_loss[k]) cudaEventDestroy(t->ev_loss[k]);
    delete t;
    return GSB_OK;
}

int gsb_trainer_bind_model(gsb_trainer* t, int count, int capacity, int shDegree, int shCoeffs, float* devLocations,
                           float* devShs, float* devScales, float* devOpacities, float* devRotations) {
    if (!t || count < 0 || shCoeffs <= 0) return gsb_fail_msg("gsb_trainer_bind_model: invalid argument");
    t->count = count; t->capacity = capacity; t->D = shDegree; t->M = shCoeffs;
    t->loc = devLocations; t->shs = devShs; t->scales = devScales; t->opac = devOpacities; t->rot = devRotations;
    t->plan_valid = false;
    return GSB_OK;
}

int gsb_trainer_set_truths(gsb_trainer* t, int n, const uint32_t* const* devW, const uint32_t* const* devB,
                           const gsb_camera* cameras, int total_cameras) {
    if (!t || n < 0 || (n > 0 && (!devW || !devB || !cameras))) return gsb_fail_msg("gsb_trainer_set_truths: invalid argument");
    t->truthW.assign(devW, devW + n);
    t->truthB.assign(devB, devB + n);
    // Trainer::train hands the same cameras over on every iteration (compat/Trainer.cpp, app_driver.cpp): the device
    // copy and its synchronisation are skipped when nothing changed
    const bool same_cams = t->d_cams && (size_t)n == t->cams.size() && (size_t)n <= t->d_cams_cap &&
                           (n == 0 || memcmp(t->cams.data(), cameras, (size_t)n * sizeof(gsb_camera)) == 0);
    t->cams.assign(cameras, cameras + n);
    t->total_cameras = total_cameras > 0 ? total_cameras : n;
    if ((size_t)n > t->d_cams_cap) {
        if (t->d_cams) GSB_CUDA_OK(cudaFree(t->d_cams));
        GSB_CUDA_OK(cudaMalloc(&t->d_cams, (size_t)n * sizeof(gsb_camera)));
        t->d_cams_cap = (size_t)n;
    }
    if (n > 0 && !same_cams) {
        GSB_CUDA_OK(cudaMemcpyAsync(t->d_cams, cameras, (size_t)n * sizeof(gsb_camera), cudaMemcpyHostToDevice, t->stream));
        GSB_CUDA_OK(cudaStreamSynchronize(t->stream));   // `cameras` may be a temporary
    }
    if ((size_t)2 * n > t->h_R_cap) {
        if (t->h_R) GSB_CUDA_OK(cudaFreeHost(t->h_R));
        GSB_CUDA_OK(cudaHostAlloc(&t->h_R, (size_t)2 * n * sizeof(uint32_t), cudaHostAllocMapped));
        t->h_R_cap = (size_t)2 * n;
    }
    return GSB_OK;
}

int gsb_trainer_set_grad_buffer(gsb_trainer* t, float* dev_avg) {
    if (!t) return gsb_fail_msg("gsb_trainer_set_grad_buffer: null trainer");
    t->avg_ext = dev_avg;
    return GSB_OK;
}
size_t gsb_trainer_grad_floats(const gsb_trainer* t) { return t ? grad_floats(t) : 0; }
float* gsb_trainer_grad_buffer(gsb_trainer* t) { return t ? avg_ptr(t) : nullptr; }

int gsb_trainer_set_allreduce(gsb_trainer* t, gsb_allreduce_fn fn, void* user) {
    if (!t) return gsb_fail_msg("gsb_trainer_set_allreduce: null trainer");
    t->allreduce = fn; t->allreduce_user = user;
    return GSB_OK;
}


int gsb_trainer_set_comm(gsb_trainer* t, gsb_comm* comm) {
    if (!t) return gsb_fail_msg("gsb_trainer_set_comm: null trainer");
    if (comm == t->comm) return GSB_OK;
    if (t->win) {                      // collective, like the call that created it
        GSB_CUDA_OK(cudaDeviceSynchronize());
        if (int e = drop_window(t)) return e;
    }
    t->comm = comm;
    return GSB_OK;
}

int gsb_trainer_exchange_path(const gsb_trainer* t) {
    if (!t || !t->comm || gsb_comm_world(t->comm) <= 1) return 0;
    return (t->win && !t->avg_ext && gsb_window_peer_mapped(t->win)) ? 2 : 1;
}

int gsb_trainer_set_fuse_backgrounds(gsb_trainer* t, int on) {
    if (!t) return gsb_fail_msg("gsb_trainer_set_fuse_backgrounds: null trainer");
    t->fuse_wb = on != 0;
    return GSB_OK;
}

int gsb_trainer_enable_loss(gsb_trainer* t, int on) {
    if (!t) return gsb_fail_msg("gsb_trainer_enable_loss: null trainer");
    t->loss_on = on != 0;
    return GSB_OK;
}

int gsb_trainer_get_stats(gsb_trainer* t, gsb_trainer_stats* out) {
    if (!t || !out) return gsb_fail_msg("gsb_trainer_get_stats: invalid argument");
    if (t->loss_pending) {
        GSB_CUDA_OK(cudaEventSynchronize(t->ev_loss[t->loss_iter & 1]));
        t->stats.sum_sq_residual = loss_of_slot(t, t->loss_iter & 1);
        t->loss_pending = false;
    }
    *out = t->stats;
    return GSB_OK;
}

int gsb_trainer_get_loss(gsb_trainer* t, int iterations_back, double* out) {
    if (!t || !out || iterations_back < 0 || iterations_back > 1) return gsb_fail_msg("gsb_trainer_get_loss: invalid argument");
    if (!t->loss_on || t->loss_iter < (unsigned)iterations_back + 1u) { gsb_fail_msg("gsb_trainer_get_loss: no such iteration (enable the loss first)"); return GSB_ERR_STATE; }
    const unsigned slot = (t->loss_iter - (unsigned)iterations_back) & 1u;
    GSB_CUDA_OK(cudaEventSynchronize(t->ev_loss[slot]));
    *out = loss_of_slot(t, slot);
    return GSB_OK;
}

int gsb_trainer_accumulate(gsb_trainer* t) {
    if (!t) return gsb_fail_msg("gsb_trainer_accumulate: null trainer");
    if (t->truthW.empty()) {
        gsb_fail_msg("Can't run training iteration, no truth data available!");   // src/Trainer.cu:253
        return GSB_ERR_STATE;
    }
    if (t->count < 0 || (t->count > 0 && !t->loc)) { gsb_fail_msg("gsb_trainer_accumulate: no model bound"); return GSB_ERR_STATE; }
    cudaStream_t s = t->stream;
    const int n = (int)t->truthW.size();
    if (t->count == 0) {
        // the reference's temporary 0-splat model (src/Trainer.cu:112) or a densify that culled everything: every
        // kernel of the iteration runs over zero splats, i.e. nothing happens
        t->stats.retries = 0;
        t->stats.rendered_total = 0;
        t->stats.views = 2 * n;
        t->stats.sum_sq_residual = 0.0;
        return GSB_OK;
    }
    const int P = t->count;
    if (int e = ensure_geom(t, (size_t)P)) return e;
    if (int e = ensure_avg(t)) return e;
    if (t->r_cap == 0) {
        if (int e = ensure_bin(t, (uint32_t)std::min<size_t>(0x7FFFFFF0u, std::max<size_t>((size_t)1 << 20, (size_t)P * 6)))) return e;
    }
    float* avg = avg_ptr(t);
    const float S = 2.0f * (float)t->total_cameras;
    t->stats.retries = 0;

    if (int e = clean_if_dirty(t)) return e;
    for (int attempt = 0; attempt < 4; attempt++) {
        GSB_NVTX("gsb accumulate");
        const bool fused = t->fuse_wb;
        const int passes = fused ? n : 2 * n;
        if (t->timing_on) {
            while (t->ev_pool.size() < (size_t)passes * (ST_COUNT + 1)) {
                cudaEvent_t e;
                GSB_CUDA_OK(cudaEventCreate(&e));
                t->ev_pool.push_back(e);
            }
        }
        // Pipelined mode: forward chains on sF, backward chains on sB, alternating workspace slots; a
        // slot is reused only after the backward chain that read it has finished.  With stage timing on
        // everything runs on `stream`, one kernel at a time, so a stage's time is its own.
        const bool pipe = t->overlap && !t->timing_on && passes > 1;
        if (pipe) {
            GSB_CUDA_OK(cudaEventRecord(t->ev_fork, s));
            GSB_CUDA_OK(cudaStreamWaitEvent(t->sF, t->ev_fork, 0));
            GSB_CUDA_OK(cudaStreamWaitEvent(t->sF2, t->ev_fork, 0));
            GSB_CUDA_OK(cudaStreamWaitEvent(t->sB, t->ev_fork, 0));
            t->exec_f = t->sF; t->exec_b = t->sB;
        }
        // zeroing the running means is only needed by the first preprocess backward: on the backward stream it runs
        // underneath the first forward chain instead of in front of it
        cudaStream_t sz = pipe ? t->sB : s;
        GSB_CUDA_OK(cudaMemsetAsync(avg, 0, grad_floats(t) * sizeof(float), sz));
        if (t->loss_on) GSB_CUDA_OK(cudaMemsetAsync(t->d_loss, 0, GSB_LOSS_SLOTS * sizeof(double), sz));
        int err = GSB_OK;
        for (int i = 0; i < passes && !err; i++) {
            if (pipe) {
                activate(t, i & 1);
                // the two slots' forward chains are independent of each other: on two streams the first two chains of
                // an iteration (nothing else to overlap with yet) share the machine instead of running back to back
                t->exec_f = (t->two_fwd_streams && (i & 1)) ? t->sF2 : t->sF;
                if (i >= 2) err = cudaStreamWaitEvent(t->exec_f, t->ev_bdone[i & 1], 0) == cudaSuccess ? GSB_OK : gsb_fail_msg("cudaStreamWaitEvent failed");
            }
            if (!err) err = view_pass(t, i, fused, avg, S, t->timing_on ? &t->ev_pool[(size_t)i * (ST_COUNT + 1)] : nullptr,
                                      i == passes - 1 ? t->ev_scan : (i == passes - 2 ? t->ev_scan_prev : nullptr));
        }
        if (pipe) {
            t->exec_f = t->exec_b = nullptr;
            cudaEventRecord(t->ev_join, t->sB);          // sB's last chain waited on sF's last chain
            cudaStreamWaitEvent(s, t->ev_join, 0);
            activate(t, 0);
        }
        if (err) { t->dirty = true; return err; }
        GSB_CUDA_OK(cudaGetLastError());
        // The host needs every view's R (arena overflow check) before anything may touch the parameters.  It waits for
        // the LAST TILE SCAN of each forward stream only (the scans of all earlier passes precede them): the last render
        // forward / backward chain is still queued behind it, so the GPU stays busy while the host checks, returns and
        // enqueues the all-reduce, apply and the next iteration.  Stage timing needs the whole iteration finished.
        if (t->timing_on) GSB_CUDA_OK(cudaStreamSynchronize(s));
        else {
            if (pipe && passes >= 2) GSB_CUDA_OK(cudaEventSynchronize(t->ev_scan_prev));
            GSB_CUDA_OK(cudaEventSynchronize(t->ev_scan));
        }
        uint32_t maxR = 0;
        long long total = 0;
        if (fused)
            for (int i = 0; i < n; i++) t->h_R[n + i] = t->h_R[i];   // the black pass shares the white pass's lists
        for (int i = 0; i < 2 * n; i++) { maxR = std::max(maxR, t->h_R[i]); total += t->h_R[i]; }
        t->stats.rendered_total = total;
        t->stats.views = 2 * n;
        if (maxR <= t->r_cap) {
            if (t->timing_on) {
                for (int i = 0; i < passes; i++)
                    for (int k = 0; k < ST_COUNT; k++) {
                        float ms = 0.f;
                        const cudaEvent_t* ev = &t->ev_pool[(size_t)i * (ST_COUNT + 1)];
                        GSB_CUDA_OK(cudaEventElapsedTime(&ms, ev[k], ev[k + 1]));
                        t->stage_ms[k] += ms;
                    }
                t->stage_passes += passes;
            }
            if (t->loss_on) {   // read back lazily: gsb_trainer_get_stats waits for this copy, train() does not
                t->loss_iter++;
                GSB_CUDA_OK(cudaMemcpyAsync(t->h_loss + (size_t)(t->loss_iter & 1) * GSB_LOSS_SLOTS, t->d_loss, GSB_LOSS_SLOTS * sizeof(double), cudaMemcpyDeviceToHost, s));
                GSB_CUDA_OK(cudaEventRecord(t->ev_loss[t->loss_iter & 1], s));
                t->loss_pending = true;
            }
            return GSB_OK;
        }
        // Some view did not fit: its gradients are incomplete.  Nothing but the running means and the
        // (self-clearing) scratch was written, so grow the arena and redo the iteration.
        t->stats.retries++;
        GSB_CUDA_OK(cudaStreamSynchronize(s));
        save_slot(t);
        for (int k = 0; k < 2; k++) {
            GSB_CUDA_OK(cudaMemsetAsync(t->slot[k].g.grads, 0, (size_t)t->geom_P * sizeof(GradRec), s));
            GSB_CUDA_OK(cudaMemsetAsync(t->slot[k].grads2, 0, (size_t)t->geom_P * sizeof(GradRec2), s));
        }
        if (int e = ensure_bin(t, (uint32_t)std::min<size_t>(0x7FFFFFF0u, (size_t)maxR + maxR / 4 + 4096))) return e;
    }
    gsb_fail_msg("gsb_trainer_accumulate: binning arena kept overflowing");
    return GSB_ERR_STATE;
}

int gsb_trainer_profile_pass(gsb_trainer* t, int pass_index, float* stage_ms, unsigned int* counters) {
    if (!t || !stage_ms) return gsb_fail_msg("gsb_trainer_profile_pass: invalid argument");
    const int n = (int)t->truthW.size();
    const bool fused = t->fuse_wb;
    if (n == 0 || pass_index < 0 || pass_index >= (fused ? n : 2 * n) || t->count <= 0 || !t->geom || !t->bin) {
        gsb_fail_msg("gsb_trainer_profile_pass: run gsb_trainer_accumulate once first");
        return GSB_ERR_STATE;
    }
    cudaEvent_t ev[ST_COUNT + 1];
    for (int k = 0; k <= ST_COUNT; k++) GSB_CUDA_OK(cudaEventCreate(&ev[k]));
    if (int e = clean_if_dirty(t)) return e;
    const int e = view_pass(t, pass_index, fused, avg_ptr(t), 2.0f * (float)t->total_cameras, ev);
    if (e) { t->dirty = true; return e; }
    GSB_CUDA_OK(cudaStreamSynchronize(t->stream));
    for (int k = 0; k < ST_COUNT; k++) GSB_CUDA_OK(cudaEventElapsedTime(&stage_ms[k], ev[k], ev[k + 1]));
    for (int k = 0; k <= ST_COUNT; k++) cudaEventDestroy(ev[k]);
    if (counters) {
        counters[0] = t->h_R[pass_index];
        GSB_CUDA_OK(cudaMemcpy(&counters[1], &t->g.ctrl->num_visible, sizeof(unsigned int), cudaMemcpyDeviceToHost));
    }
    return GSB_OK;
}

int gsb_trainer_debug_set_epoch(gsb_trainer* t, unsigned int epoch) {
    if (!t) return gsb_fail_msg("gsb_trainer_debug_set_epoch: null trainer");
    t->epoch = epoch;
    return GSB_OK;
}

int gsb_trainer_set_overlap(gsb_trainer* t, int on) {
    if (!t) return gsb_fail_msg("gsb_trainer_set_overlap: null trainer");
    t->overlap = on != 0;
    return GSB_OK;
}

int gsb_trainer_set_stage_timing(gsb_trainer* t, int on) {
    if (!t) return gsb_fail_msg("gsb_trainer_set_stage_timing: null trainer");
    t->timing_on = on != 0;
    for (int k = 0; k < GSB_NUM_STAGES; k++) t->stage_ms[k] = 0.0;
    t->stage_passes = 0;
    t->stats.exchange_ms = t->stats.apply_ms = 0.0;
    t->stats.timed_iterations = 0;
    return GSB_OK;
}

int gsb_trainer_get_stage_times(gsb_trainer* t, double* stage_ms, long long* passes) {
    if (!t || !stage_ms || !passes) return gsb_fail_msg("gsb_trainer_get_stage_times: invalid argument");
    for (int k = 0; k < GSB_NUM_STAGES; k++) stage_ms[k] = t->stage_ms[k];
    *passes = t->stage_passes;
    return GSB_OK;
}

int gsb_trainer_count_pairs(gsb_trainer* t, int camera_index, unsigned long long* out8) {
    if (!t || !out8) return gsb_fail_msg("gsb_trainer_count_pairs: invalid argument");
    const int n = (int)t->truthW.size();
    if (n == 0 || camera_index < 0 || camera_index >= n || t->count <= 0 || !t->geom || !t->bin) {
        gsb_fail_msg("gsb_trainer_count_pairs: run gsb_trainer_accumulate once first");
        return GSB_ERR_STATE;
    }
    cudaStream_t s = t->stream;
    if (!t->d_pairs) GSB_CUDA_OK(cudaMalloc(&t->d_pairs, 6 * sizeof(unsigned long long)));
    GSB_CUDA_OK(cudaMemsetAsync(t->d_pairs, 0, 6 * sizeof(unsigned long long), s));
    const gsb_camera* dc = t->d_cams + camera_index;
    if (int e = clean_if_dirty(t)) return e;
    if (int e = forward_pass(t, dc->view, dc->projview, dc->location, t->cams[camera_index].tan_fovx,
                             t->cams[camera_index].tan_fovy, 1.0f, t->d_bg + 3, t->W, t->H, t->im, t->rendered, t->h_R)) {
        t->dirty = true;
        return e;
    }
    RenderFwdArgs ra;
    ra.W = t->W; ra.H = t->H; ra.grid_x = t->grid_x; ra.grid_y = t->grid_y;
    ra.ranges = t->im.ranges; ra.point_list = t->b.point_list; ra.recs = t->g.recs; ra.bg = t->d_bg + 3;
    ra.out_color = t->rendered; ra.final_T = t->im.final_T; ra.n_contrib = t->im.n_contrib; ra.tile_max = t->im.tile_max;
    ra.frame = nullptr; ra.tile_order = nullptr; ra.ctrl = nullptr;
    launch_count_pairs(ra, t->d_pairs, s);
    GSB_CUDA_OK(cudaMemcpyAsync(out8, t->d_pairs, 6 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    uint32_t cnt[2] = {0, 0};
    GSB_CUDA_OK(cudaMemcpyAsync(cnt, &t->g.ctrl->num_rendered, sizeof(cnt), cudaMemcpyDeviceToHost, s));
    GSB_CUDA_OK(cudaStreamSynchronize(s));
    out8[6] = cnt[0];
    out8[7] = cnt[1];
    return GSB_OK;
}

int gsb_trainer_measure_bwd_skeleton(gsb_trainer* t, int camera_index, int reps, float* out_ms2) {
    if (!t || !out_ms2 || reps <= 0) return gsb_fail_msg("gsb_trainer_measure_bwd_skeleton: invalid argument");
    const int n = (int)t->truthW.size();
    if (n == 0 || camera_index < 0 || camera_index >= n || t->count <= 0 || !t->geom || !t->bin) {
        gsb_fail_msg("gsb_trainer_measure_bwd_skeleton: run gsb_trainer_accumulate once first");
        return GSB_ERR_STATE;
    }
    cudaStream_t s = t->stream;
    if (int e = clean_if_dirty(t)) return e;
    const gsb_camera* dc = t->d_cams + camera_index;
    if (int e = forward_pass(t, dc->view, dc->projview, dc->location, t->cams[camera_index].tan_fovx,
                             t->cams[camera_index].tan_fovy, 1.0f, t->d_bg + 3, t->W, t->H, t->im, t->rendered, t->h_R)) {
        t->dirty = true;
        return e;
    }
    RenderBwdArgs rb;
    rb.W = t->W; rb.H = t->H; rb.grid_x = t->grid_x; rb.grid_y = t->grid_y;
    rb.ranges = t->im.ranges; rb.point_list = t->b.point_list; rb.recs = t->g.recs; rb.bg = t->d_bg + 3;
    rb.final_T = t->im.final_T; rb.n_contrib = t->im.n_contrib; rb.tile_max = t->im.tile_max;
    rb.dL_dpix = nullptr; rb.truth = t->truthW[camera_index]; rb.rendered = t->rendered; rb.grads = t->g.grads;
    rb.truth2 = t->truthB[camera_index]; rb.grads2 = t->grads2; rb.loss_out = nullptr;
    rb.tile_order = t->im.tile_order; rb.ctrl = t->g.ctrl;
    cudaEvent_t e0, e1;
    GSB_CUDA_OK(cudaEventCreate(&e0));
    GSB_CUDA_OK(cudaEventCreate(&e1));
    for (int mode = 0; mode < 2; mode++) {      // 0: the shipped render_bwd_kernel<2>, 1: the lanes = splats skeleton
        float best = 1e30f;
        for (int r = 0; r < reps + 1; r++) {
            GSB_CUDA_OK(cudaEventRecord(e0, s));
            if (mode == 0) launch_render_bwd(rb, s); else launch_bwd_skeleton(rb, s);
            GSB_CUDA_OK(cudaEventRecord(e1, s));
            GSB_CUDA_OK(cudaEventSynchronize(e1));
            float ms = 0.f;
            GSB_CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
            if (r > 0) best = std::min(best, ms);
        }
        out_ms2[mode] = best;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    GSB_CUDA_OK(cudaMemsetAsync(t->grads2, 0, (size_t)t->geom_P * sizeof(GradRec2), s));   // the scratch must be clean again
    GSB_CUDA_OK(cudaStreamSynchronize(s));
    GSB_CUDA_OK(cudaGetLastError());
    return GSB_OK;
}

// A/B of the record staging (csrc/measure.cu staging_ab_kernel): out_ms3 = {cp.async, gather4 by lanes 0-7, gather4 by
// lane 0}, best of `reps` launches each over the tile lists of training camera `camera_index`; a mode whose copies did not
// complete reports -1.
int gsb_trainer_measure_staging(gsb_trainer* t, int camera_index, int reps, float* out_ms3) {
    if (!t || !out_ms3 || reps <= 0) return gsb_fail_msg("gsb_trainer_measure_staging: invalid argument");
    const int n = (int)t->truthW.size();
    if (n == 0 || camera_index < 0 || camera_index >= n || t->count <= 0 || !t->geom || !t->bin) {
        gsb_fail_msg("gsb_trainer_measure_staging: run gsb_trainer_accumulate once first");
        return GSB_ERR_STATE;
    }
    cudaStream_t s = t->stream;
    if (int e = clean_if_dirty(t)) return e;
    const gsb_camera* dc = t->d_cams + camera_index;
    if (int e = forward_pass(t, dc->view, dc->projview, dc->location, t->cams[camera_index].tan_fovx,
                             t->cams[camera_index].tan_fovy, 1.0f, t->d_bg + 3, t->W, t->H, t->im, t->rendered, t->h_R)) {
        t->dirty = true;
        return e;
    }
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    GSB_CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (!fn) return gsb_fail_msg("gsb_trainer_measure_staging: cuTensorMapEncodeTiled unavailable");
    CUtensorMap tm;
    const cuuint64_t dims[2] = {12, (cuuint64_t)t->count};
    const cuuint64_t strides[1] = {sizeof(SplatRec)};
    const cuuint32_t box[2] = {12, 1}, estr[2] = {1, 1};
    if (((EncodeFn)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, t->g.recs, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return gsb_fail_msg("gsb_trainer_measure_staging: cuTensorMapEncodeTiled failed");
    RenderFwdArgs ra;
    ra.W = t->W; ra.H = t->H; ra.grid_x = t->grid_x; ra.grid_y = t->grid_y;
    ra.ranges = t->im.ranges; ra.point_list = t->b.point_list; ra.recs = t->g.recs; ra.bg = t->d_bg + 3;
    ra.out_color = t->rendered; ra.final_T = t->im.final_T; ra.n_contrib = t->im.n_contrib; ra.tile_max = t->im.tile_max;
    ra.frame = nullptr; ra.tile_order = nullptr; ra.ctrl = nullptr;
    float* sink = nullptr;
    GSB_CUDA_OK(cudaMalloc(&sink, 2 * sizeof(float)));
    cudaEvent_t e0, e1;
    GSB_CUDA_OK(cudaEventCreate(&e0));
    GSB_CUDA_OK(cudaEventCreate(&e1));
    for (int mode = 0; mode < 3; mode++) {
        float best = 1e30f;
        GSB_CUDA_OK(cudaMemsetAsync(sink, 0, 2 * sizeof(float), s));
        for (int r = 0; r < reps + 1; r++) {
            GSB_CUDA_OK(cudaEventRecord(e0, s));
            launch_staging_ab(ra, &tm, mode, sink, s);
            GSB_CUDA_OK(cudaEventRecord(e1, s));
            GSB_CUDA_OK(cudaEventSynchronize(e1));
            float ms = 0.f;
            GSB_CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
            if (r > 0) best = std::min(best, ms);
        }
        float flag[2] = {0.f, 0.f};
        GSB_CUDA_OK(cudaMemcpy(flag, sink, sizeof(flag), cudaMemcpyDeviceToHost));
        out_ms3[mode] = flag[1] < 0.f ? -1.0f : best;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
    GSB_CUDA_OK(cudaGetLastError());
    return GSB_OK;
}

// Statistics of the render backward's reductions for training camera `camera_index` (csrc/measure.cu
// reduction_stats_kernel): out40 as documented there.
int gsb_trainer_reduction_stats(gsb_trainer* t, int camera_index, unsigned long long* out40) {
    if (!t || !out40) return gsb_fail_msg("gsb_trainer_reduction_stats: invalid argument");
    const int n = (int)t->truthW.size();
    if (n == 0 || camera_index < 0 || camera_index >= n || t->count <= 0 || !t->geom || !t->bin) {
        gsb_fail_msg("gsb_trainer_reduction_stats: run gsb_trainer_accumulate once first");
        return GSB_ERR_STATE;
    }
    cudaStream_t s = t->stream;
    if (int e = clean_if_dirty(t)) return e;
    const gsb_camera* dc = t->d_cams + camera_index;
    if (int e = forward_pass(t, dc->view, dc->projview, dc->location, t->cams[camera_index].tan_fovx,
                             t->cams[camera_index].tan_fovy, 1.0f, t->d_bg + 3, t->W, t->H, t->im, t->rendered, t->h_R)) {
        t->dirty = true;
        return e;
    }
    RenderBwdArgs rb;
    memset(&rb, 0, sizeof(rb));
    rb.W = t->W; rb.H = t->H; rb.grid_x = t->grid_x; rb.grid_y = t->grid_y;
    rb.ranges = t->im.ranges; rb.point_list = t->b.point_list; rb.recs = t->g.recs; rb.bg = t->d_bg + 3;
    rb.final_T = t->im.final_T; rb.n_contrib = t->im.n_contrib; rb.tile_max = t->im.tile_max;
    unsigned long long* d = nullptr;
    GSB_CUDA_OK(cudaMalloc(&d, 40 * sizeof(unsigned long long)));
    GSB_CUDA_OK(cudaMemsetAsync(d, 0, 40 * sizeof(unsigned long long), s));
    launch_reduction_stats(rb, d, s);
    GSB_CUDA_OK(cudaMemcpyAsync(out40, d, 40 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    GSB_CUDA_OK(cudaStreamSynchronize(s));
    cudaFree(d);
    GSB_CUDA_OK(cudaGetLastError());
    return GSB_OK;
}

int gsb_trainer_apply(gsb_trainer* t, const gsb_project_params* p) {
    if (!t || !p) return gsb_fail_msg("gsb_trainer_apply: invalid argument");
    ApplyArgs a;
    a.locations = t->loc; a.shs = t->shs; a.scales = t->scales; a.opacities = t->opac; a.rotations = t->rot;
    a.avg = avg_ptr(t);
    a.lrLocation = p->lrLocation; a.lrSh = p->lrSh; a.lrScale = p->lrScale; a.lrOpacity = p->lrOpacity;
    a.lrRotation = p->lrRotation; a.maxScale = p->paramScaleMax; a.M = t->M; a.count = t->count;
    launch_apply_gradients(a, t->stream);
    t->stats.kernel_launches++;
    GSB_CUDA_OK(cudaGetLastError());
    return GSB_OK;
}

int gsb_trainer_densify_plan(gsb_trainer* t, const gsb_project_params* p, int* new_count) {
    if (!t || !p || !new_count) return gsb_fail_msg("gsb_trainer_densify_plan: invalid argument");
    cudaStream_t s = t->stream;
    const size_t P = (size_t)t->count;
    if (P > t->dens_P) {
        cudaFree(t->flags); cudaFree(t->keep_idx); cudaFree(t->aux_idx); cudaFree(t->dens_lb);
        const size_t cap = P + P / 4 + 1024;
        GSB_CUDA_OK(cudaMalloc(&t->flags, cap));
        GSB_CUDA_OK(cudaMalloc(&t->keep_idx, cap * sizeof(uint32_t)));
        GSB_CUDA_OK(cudaMalloc(&t->aux_idx, cap * sizeof(uint32_t)));
        const size_t lbw = ((cap + 1023) / 1024 + 1) * 4;
        GSB_CUDA_OK(cudaMalloc(&t->dens_lb, lbw * sizeof(unsigned long long)));
        GSB_CUDA_OK(cudaMemsetAsync(t->dens_lb, 0, lbw * sizeof(unsigned long long), s));
        t->dens_P = cap;
    }
    if (int e = epoch_guard(t)) return e;
    GSB_CUDA_OK(cudaMemsetAsync(t->dens_ticket, 0, sizeof(uint32_t), s));
    GSB_CUDA_OK(cudaMemsetAsync(t->totals, 0, 4 * sizeof(uint32_t), s));
    DensifyArgs a;
    a.locations = t->loc; a.shs = t->shs; a.scales = t->scales; a.opacities = t->opac; a.rotations = t->rot;
    a.avg = avg_ptr(t); a.M = t->M; a.count = t->count;
    a.cullOpacity = p->paramCullOpacity; a.cullSize = p->paramCullSize; a.densifyVariance = p->paramDensifyVariance;
    a.splitSize = p->paramSplitSize; a.splitDistance = p->paramSplitDistance; a.splitScale = p->paramSplitScale;
    a.cloneDistance = p->paramCloneDistance;
    a.flags = t->flags; a.keep_idx = t->keep_idx; a.aux_idx = t->aux_idx; a.totals = t->totals;
    a.lookback = t->dens_lb; a.epoch = ++t->epoch; a.ticket = t->dens_ticket;
    launch_densify_classify_scan(a, s);
    t->stats.kernel_launches++;
    uint32_t tot[3] = {0, 0, 0};
    GSB_CUDA_OK(cudaMemcpyAsync(tot, t->totals, sizeof(tot), cudaMemcpyDeviceToHost, s));
    GSB_CUDA_OK(cudaStreamSynchronize(s));
    // The reference appends while count < capacity, counting before the cull compaction
    // (src/Trainer.cu:460,500): splits first, then clones.
    const long long room = std::max(0LL, (long long)t->capacity - (long long)t->count);
    t->n_keep = tot[0];
    t->n_split_done = (uint32_t)std::min<long long>(tot[1], room);
    t->n_clone_done = (uint32_t)std::min<long long>(tot[2], room - t->n_split_done);
    t->plan_params = *p;
    t->plan_valid = true;
    *new_count = (int)(t->n_keep + t->n_split_done + t->n_clone_done);
    return GSB_OK;
}

int gsb_trainer_densify_commit(gsb_trainer* t, float* devLocations, float* devShs, float* devScales, float* devOpacities,
                               float* devRotations) {
    if (!t || !t->plan_valid) { gsb_fail_msg("gsb_trainer_densify_commit: no plan"); return GSB_ERR_STATE; }
    const gsb_project_params* p = &t->plan_params;
    DensifyArgs a;
    a.locations = t->loc; a.shs = t->shs; a.scales = t->scales; a.opacities = t->opac; a.rotations = t->rot;
    a.avg = avg_ptr(t); a.M = t->M; a.count = t->count;
    a.cullOpacity = p->paramCullOpacity; a.cullSize = p->paramCullSize; a.densifyVariance = p->paramDensifyVariance;
    a.splitSize = p->paramSplitSize; a.splitDistance = p->paramSplitDistance; a.splitScale = p->paramSplitScale;
    a.cloneDistance = p->paramCloneDistance;
    a.flags = t->flags; a.keep_idx = t->keep_idx; a.aux_idx = t->aux_idx; a.totals = t->totals;
    a.lookback = t->dens_lb; a.epoch = t->epoch; a.ticket = t->dens_ticket;
    DensifyOut o;
    o.locations = devLocations; o.shs = devShs; o.scales = devScales; o.opacities = devOpacities; o.rotations = devRotations;
    o.n_keep = t->n_keep; o.n_split_done = t->n_split_done; o.n_clone_done = t->n_clone_done;
    launch_densify_scatter(a, o, t->stream);
    t->stats.kernel_launches++;
    GSB_CUDA_OK(cudaGetLastError());
    GSB_CUDA_OK(cudaStreamSynchronize(t->stream));   // the caller frees the old arrays next
    t->count = (int)(t->n_keep + t->n_split_done + t->n_clone_done);
    t->loc = devLocations; t->shs = devShs; t->scales = devScales; t->opac = devOpacities; t->rot = devRotations;
    t->plan_valid = false;
    return GSB_OK;
}

int gsb_trainer_train(gsb_trainer* t, const gsb_project_params* p, int densify, int* new_count) {
    if (!t || !p) return gsb_fail_msg("gsb_trainer_train: invalid argument");
    GSB_NVTX("gsb train");
    if (int e = gsb_trainer_accumulate(t)) return e;
    // with stage timing on, the exchange and the update are bracketed too (events on the trainer's stream)
    const bool timed = t->timing_on;
    if (timed && !t->ev_x[0])
        for (int k = 0; k < 3; k++) GSB_CUDA_OK(cudaEventCreate(&t->ev_x[k]));
    if (timed) GSB_CUDA_OK(cudaEventRecord(t->ev_x[0], t->stream));
    if (t->comm) {
        GSB_NVTX("exchange");
        if (t->win && !t->avg_ext) {
            if (int e = gsb_window_allreduce_sum(t->win, grad_floats(t), t->stream)) return e;    // peer kernel (or NCCL inside)
        } else if (int e = gsb_comm_allreduce_sum(t->comm, avg_ptr(t), grad_floats(t), t->stream)) return e;
    } else if (t->allreduce) {
        GSB_NVTX("all-reduce hook");
        if (int e = t->allreduce(avg_ptr(t), grad_floats(t), t->stream, t->allreduce_user)) {
            gsb_fail_msg("gsb_trainer_train: all-reduce hook failed");
            return e;
        }
    }
    if (timed) GSB_CUDA_OK(cudaEventRecord(t->ev_x[1], t->stream));
    if (int e = gsb_trainer_apply(t, p)) return e;
    if (timed) {
        GSB_CUDA_OK(cudaEventRecord(t->ev_x[2], t->stream));
        GSB_CUDA_OK(cudaEventSynchronize(t->ev_x[2]));
        float a = 0.f, b = 0.f;
        GSB_CUDA_OK(cudaEventElapsedTime(&a, t->ev_x[0], t->ev_x[1]));
        GSB_CUDA_OK(cudaEventElapsedTime(&b, t->ev_x[1], t->ev_x[2]));
        t->stats.exchange_ms += a; t->stats.apply_ms += b; t->stats.timed_iterations++;
    }
    if (densify) {
        if (!new_count) return gsb_fail_msg("gsb_trainer_train: densify needs new_count");
        return gsb_trainer_densify_plan(t, p, new_count);
    }
    return GSB_OK;
}

int gsb_trainer_render(gsb_trainer* t, uint32_t* dev_frame_buffer, int width, int height, float splat_scale,
                       const gsb_camera* camera) {
    if (!t || !dev_frame_buffer || !camera || width <= 0 || height <= 0) return gsb_fail_msg("gsb_trainer_render: invalid argument");
    cudaStream_t s = t->stream;
    const size_t N = (size_t)width * height;
    const int tiles = ((width + GSB_TILE - 1) / GSB_TILE) * ((height + GSB_TILE - 1) / GSB_TILE);
    if (int e = ensure_geom(t, (size_t)std::max(t->count, 1))) return e;
    if (t->r_cap == 0) {
        if (int e = ensure_bin(t, (uint32_t)std::min<size_t>(0x7FFFFFF0u, std::max<size_t>((size_t)1 << 20, (size_t)t->count * 6)))) return e;
    }
    // preview arena: the training image workspace is reused when the size matches (T:159-167)
    if (int e = clean_if_dirty(t)) return e;
    ImageWs im;
    float* color;
    if (width == t->W && height == t->H) {
        im = t->im; color = t->rendered;
    } else {
        const size_t need = ImageWs::required(N, (size_t)tiles);
        if (need > t->pimg_bytes) {
            GSB_CUDA_OK(cudaStreamSynchronize(s));
            cudaFree(t->pimg);
            GSB_CUDA_OK(cudaMalloc(&t->pimg, need));
            t->pimg_bytes = need;
            t->pimg_N = 0;
        }
        if (3 * N > t->pcolor_floats) {
            GSB_CUDA_OK(cudaStreamSynchronize(s));
            cudaFree(t->pcolor);
            GSB_CUDA_OK(cudaMalloc(&t->pcolor, 3 * N * sizeof(float)));
            t->pcolor_floats = 3 * N;
        }
        if (t->pimg_N != N || t->pimg_tiles != (size_t)tiles) {   // new preview size: re-carve, housekeeping arrays start at zero
            t->pim.carve(t->pimg, N, (size_t)tiles);
            t->pimg_N = N; t->pimg_tiles = (size_t)tiles;
            GSB_CUDA_OK(cudaMemsetAsync(t->pim.tile_max, 0, t->pim.zero_bytes((size_t)tiles), s));
        }
        im = t->pim;
        color = t->pcolor;
    }
    gsb_camera* dc = t->d_pcam;
    GSB_CUDA_OK(cudaMemcpyAsync(dc, camera, sizeof(gsb_camera), cudaMemcpyHostToDevice, s));
    for (int attempt = 0; attempt < 4; attempt++) {
        uint32_t* hR = nullptr;
        if (!t->h_R) {
            GSB_CUDA_OK(cudaHostAlloc(&t->h_R, 2 * sizeof(uint32_t), cudaHostAllocMapped));
            t->h_R_cap = 2;
        }
        hR = t->h_R;
        if (int e = forward_pass(t, dc->view, dc->projview, dc->location, camera->tan_fovx, camera->tan_fovy, splat_scale,
                                 t->d_bg + 3, width, height, im, color, hR, nullptr, nullptr, dev_frame_buffer)) {
            t->dirty = true;
            return e;
        }
        GSB_CUDA_OK(cudaStreamSynchronize(s));   // the reference ends render() with a device sync (T:215)
        if (t->count == 0 || hR[0] <= t->r_cap) break;
        if (int e = ensure_bin(t, (uint32_t)std::min<size_t>(0x7FFFFFF0u, (size_t)hR[0] + hR[0] / 4 + 4096))) return e;
    }
    return GSB_OK;
}

}  // extern "C"
