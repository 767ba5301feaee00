// api.cu -- the C-ABI of libsvtme_cuda.so (include/svtme.h): context, picture slots, batched ME.
// Host side only; every kernel lives in its own translation unit and is reached through launch_*().
//
// Three streams per context so that transfers overlap the searches:
//   copy_in  : H2D of uploaded luma planes + pyramid_kernel; records Slot::ready
//   stream   : the ME kernels of each batch (waits on the ready events of the pictures it reads)
//   copy_out : D2H of a batch's results into the caller's buffers (waits on Batch::computed)
// Batches rotate through a small ring of job tables / scratch / result buffers, so a new batch can be
// enqueued while the previous one computes and the one before that drains.
#include "me_common.cuh"
#include <cuda.h>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <algorithm>
#include <vector>

using namespace svtme;

namespace svtme {
cudaError_t configure_fullpel();
cudaError_t configure_fullpel_ext();
cudaError_t configure_subpel();
}

constexpr int kRing = 3;

// Contexts alive in this process: svtme_picture_upload_peer leaves events of the COPYING context in the slots of the
// OWNING one, so a context that is destroyed has to take its events back out of every live peer first.
#include <set>
static std::mutex g_live_mu;
static std::set<SvtMeContext*> g_live;

struct Batch {
    int cap = 0;                          // pictures this entry can hold
    SbScratch* d_scratch = nullptr;       // [cap][n_sb]
    SbScratchExt* d_scratch_ext = nullptr;// [cap][n_sb], allocated when a batch first holds a 209-PU job
    SvtMePuResult* d_results = nullptr;   // [cap][n_sb][res_pus] rows; a job's rows are packed with its own num_pus
    int res_pus = 0;                      // 85, or 209 once a batch held a 209-PU job
    size_t res_off[SVTME_MAX_BATCH];      // first result row of every job of the batch in d_results
    SvtMeSbDebug* d_debug = nullptr;      // [cap][n_sb]
    DJob* d_jobs = nullptr;               // [cap]
    uint32_t* d_worklist = nullptr;       // [cap][n_sb + 1]: deferred-SB lists of the sub-pel passes, one region per chunk (counter first)
    DJob* h_jobs = nullptr;               // pinned
    cudaEvent_t computed = nullptr;       // kernels of the last batch that used this entry are done
    cudaEvent_t drained = nullptr;        // ... and its results have reached the host
    bool used = false;
};

struct Slot {
    uint8_t* base = nullptr;      // one allocation: luma staging + three padded planes
    uint8_t* luma = nullptr;      // unpadded luma as uploaded (pitch = width)
    DPyramid pyr;
    size_t plane_bytes[3] = { 0, 0, 0 };
    cudaEvent_t ready = nullptr;  // pyramid of the last upload is complete
    cudaEvent_t ready_ipc = nullptr;   // same instant, inter-process flavour (exported by svtme_slot_export)
    int last_reader = -1;         // ring entry of the last batch that reads this slot ...
    int64_t last_reader_seq = -1; // ... and that batch's sequence number
    std::vector<cudaEvent_t> peer_reads;   // copies other contexts took of this slot since its last rewrite (owned by them)
};

struct SvtMeContext {
    int device = 0, width = 0, height = 0, sb_cols = 0, sb_rows = 0, n_sb = 0;
    cudaStream_t stream = nullptr, copy_in = nullptr, copy_out = nullptr;
    uint8_t* d_tmaps = nullptr;           // SVTME_TMA=1: one 128-byte CUtensorMap per slot (its padded full plane)
    uint8_t* d_zz = nullptr;              // svtme_zz_sad: [cap] ZzJob, then [cap][2][n_sb] output bytes
    int zz_cap = 0;
    uint8_t* d_ois = nullptr;             // svtme_ois: [cap] OisJob, then per job me_sad / cand / total
    int ois_cap = 0;
    int ois_variant = 0;                  // SVTME_OIS_KERNEL: 0 = line-per-lane kernel (default), 1 = pixel-wise, 2 = word-wise (earlier generations, kept for A/B)
    cudaEvent_t join_ev = nullptr;
    // chunked submission: a small batch (one mini-GOP: a handful of waves per kernel) is cut into two chunks of pictures that
    // run their kernel sequences on two streams, so the tail of one chunk's kernel is filled by the other chunk's CTAs.
    // SVTME_CHUNKS overrides (1 = never); large batches and profiled runs (per-kernel event times) stay on one stream.
    static constexpr int kChunkStreams = 8;
    int chunks = 0;                       // 0 = automatic
    int chunk_streams = 2;                // streams the chunks alternate on
    cudaStream_t cstream[kChunkStreams] = {};
    cudaEvent_t fork_ev = nullptr, cjoin_ev[kChunkStreams] = {};
    std::vector<Slot> slots;
    std::mutex peer_mu;                   // guards Slot::peer_reads (written by the contexts that copy from us)
    std::vector<cudaEvent_t> peer_events; // events this context records after copying from a peer (round-robin)
    size_t next_peer_event = 0;
    std::map<std::string, std::pair<void*, cudaEvent_t>> ipc_open;   // imported slots: mem handle bytes -> (base, ready event)
    uint16_t* d_in16 = nullptr;           // staging of a 16-bit luma plane (svtme_picture_upload_16bit), width * height samples
    uint8_t* d_rem = nullptr;             // its 2-bit remainder plane
    Batch ring[kRing];
    int next = 0, last = -1;              // ring entry for the next batch / of the latest batch
    int64_t seq = 0;                      // batches submitted so far
    int64_t launches = 0;
    std::string err;
    // optional per-kernel timing (CUDA events on the launching stream)
    bool profiling = false;
    struct Timed { int kind; cudaEvent_t a, b; };
    std::vector<Timed> timed;
    std::vector<cudaEvent_t> event_pool;
    double kind_ms[SVTME_KERNEL_KINDS] = {};
    int64_t kind_launches[SVTME_KERNEL_KINDS] = {};
};

// Brackets one kernel launch with events when profiling is on.
struct LaunchTimer {
    SvtMeContext* c; int kind; cudaStream_t st; cudaEvent_t a = nullptr, b = nullptr;
    static cudaEvent_t get(SvtMeContext* c) {
        cudaEvent_t e = nullptr;
        if (!c->event_pool.empty()) { e = c->event_pool.back(); c->event_pool.pop_back(); }
        else cudaEventCreate(&e);
        return e;
    }
    LaunchTimer(SvtMeContext* ctx, int k, cudaStream_t s) : c(ctx), kind(k), st(s) {
        c->launches++; c->kind_launches[k]++;
        if (c->profiling) { a = get(c); b = get(c); cudaEventRecord(a, st); }
    }
    ~LaunchTimer() {
        if (a) { cudaEventRecord(b, st); c->timed.push_back({ kind, a, b }); }
    }
};

static int32_t fail(SvtMeContext* c, int32_t code, const std::string& msg) {
    if (c) c->err = msg;
    return code;
}
#define CU(call)                                                                                   \
    do { cudaError_t e_ = (call);                                                                  \
         if (e_ != cudaSuccess)                                                                    \
             return fail(ctx, e_ == cudaErrorMemoryAllocation ? SVTME_ERR_INSUFFICIENT_RESOURCES   \
                                                                : SVTME_ERR_UNDEFINED,              \
                         std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static void describe_plane(DPlane& pl, uint8_t* mem, int w, int h, int pad) {
    pl.buf = mem; pl.stride = (int)align_up((size_t)w + 2 * pad, 64);
    pl.ox = pad; pl.oy = pad; pl.w = w; pl.h = h;
}
static size_t plane_alloc_bytes(int w, int h, int pad) {
    return align_up(align_up((size_t)w + 2 * pad, 64) * (size_t)(h + 2 * pad) + 256, 256);   // +256: tail slack for word reads
}

extern "C" {

/* EbDefinitions.h:3320-3469 column M0, EbResourceCoordinationProcess.c:366-380, EbDefinitions.h:2980-2996 */
void svtme_default_params(SvtMeParams* p, int32_t width, int32_t height, int32_t num_lists,
                          int32_t temporal_layer_index, int32_t hierarchical_levels, int32_t is_ref) {
    static const int16_t l0_total_w[5] = { 48, 64, 96, 96, 128 }, l0_w[5] = { 24, 32, 48, 48, 64 };
    static const int16_t l0_total_h[5] = { 40, 48, 48, 48, 80 }, l0_h[5] = { 20, 24, 24, 24, 40 };
    static const int8_t en_l2[5] = { 1, 1, 1, 1, 0 };
    static const int16_t mult[6][6] = { { 100 }, { 100, 100 }, { 100, 100, 100 }, { 200, 140, 100, 70 },
                                        { 350, 200, 100, 100, 100 }, { 525, 350, 200, 100, 100, 100 } };
    hierarchical_levels = std::max(0, std::min(hierarchical_levels, 5));
    temporal_layer_index = std::max(0, std::min(temporal_layer_index, hierarchical_levels));
    const uint32_t size = (uint32_t)width * (uint32_t)height;
    const int ir = size < 0xB71B0 ? 0 : size < 0x1AB3F0 ? 1 : size < 0x29F630 ? 2 : 3;
    const uint32_t ratio = (uint32_t)width / (uint32_t)height;
    const int ri = ir <= 0 ? 0 : (ir <= 1 && ratio < 2) ? 1 : (ir <= 1 && ratio > 3) ? 2 : (ir <= 2) ? 3 : 4;
    memset(p, 0, sizeof(*p));
    p->width = width; p->height = height; p->num_lists = num_lists;
    p->temporal_layer_index = temporal_layer_index; p->hierarchical_levels = hierarchical_levels;
    p->is_used_as_reference = is_ref;
    p->enable_hme_flag = 1; p->enable_hme_level0 = 1; p->enable_hme_level1 = 1; p->enable_hme_level2 = en_l2[ri];
    p->use_subpel = 1; p->num_pus = SVTME_SQUARE_PU_COUNT; p->cu8x8_mode = 0;
    p->num_hme_regions_w = 2; p->num_hme_regions_h = 2;
    p->hme_l0_total_w = l0_total_w[ri]; p->hme_l0_total_h = l0_total_h[ri];
    for (int i = 0; i < 2; i++) {
        p->hme_l0_w[i] = l0_w[ri]; p->hme_l0_h[i] = l0_h[ri];
        p->hme_l1_w[i] = 16; p->hme_l1_h[i] = 16; p->hme_l2_w[i] = 8; p->hme_l2_h[i] = 8;
    }
    p->search_area_w = 64; p->search_area_h = 64;
    p->hme_l0_mult_x = p->hme_l0_mult_y = mult[hierarchical_levels][temporal_layer_index];
}

int32_t svtme_create(SvtMeContext** out, int32_t device, int32_t width, int32_t height, int32_t num_slots) {
    if (!out) return SVTME_ERR_BAD_PARAMETER;
    *out = nullptr;
    if (width <= 0 || height <= 0 || (width & 7) || (height & 7) || num_slots <= 0 ||
        (width & 63) == 40 || (width & 63) == 56 || width > 8192 || height > 8192)   /* 40/56-wide edge SBs: void kernel in the reference, EbComputeSAD.h:143-147 */
        return SVTME_ERR_BAD_PARAMETER;
    SvtMeContext* ctx = new (std::nothrow) SvtMeContext();
    if (!ctx) return SVTME_ERR_INSUFFICIENT_RESOURCES;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        delete ctx; return SVTME_ERR_UNDEFINED;             // no GPU: fail loudly, there is no CPU path
    }
    { std::lock_guard<std::mutex> lock(g_live_mu); g_live.insert(ctx); }
    ctx->device = device; ctx->width = width; ctx->height = height;
    ctx->sb_cols = (width + 63) / 64; ctx->sb_rows = (height + 63) / 64; ctx->n_sb = ctx->sb_cols * ctx->sb_rows;
    *out = ctx;                                              // handed out even on failure so the error text is readable
    CU(cudaSetDevice(device));
    CU(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    {   // the upload stream also runs the (tiny) split / pyramid kernels: highest priority, so their CTAs are placed as soon as
        // any search CTA retires instead of waiting for the whole grid of a 10 ms search kernel to be issued
        int least = 0, greatest = 0;
        CU(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        CU(cudaStreamCreateWithPriority(&ctx->copy_in, cudaStreamNonBlocking, greatest));
    }
    CU(cudaStreamCreateWithFlags(&ctx->copy_out, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&ctx->join_ev, cudaEventDisableTiming));
    if (const char* e = getenv("SVTME_OIS_KERNEL")) ctx->ois_variant = std::max(0, std::min(atoi(e), 2));
    if (const char* e = getenv("SVTME_CHUNKS")) ctx->chunks = std::max(0, std::min(atoi(e), 64));
    if (ctx->chunks != 1) {
        CU(cudaEventCreateWithFlags(&ctx->fork_ev, cudaEventDisableTiming));
        ctx->chunk_streams = ctx->chunks ? std::min(ctx->chunks, (int)SvtMeContext::kChunkStreams) : 2;
        if (const char* e2 = getenv("SVTME_CHUNK_STREAMS")) ctx->chunk_streams = std::max(1, std::min(atoi(e2), ctx->chunk_streams));
        for (int i = 0; i < ctx->chunk_streams; i++) {
            CU(cudaStreamCreateWithFlags(&ctx->cstream[i], cudaStreamNonBlocking));
            CU(cudaEventCreateWithFlags(&ctx->cjoin_ev[i], cudaEventDisableTiming));
        }
    }
    for (auto& b : ctx->ring) {
        CU(cudaEventCreateWithFlags(&b.computed, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&b.drained, cudaEventDisableTiming));
    }
    CU(configure_fullpel());
    CU(configure_fullpel_ext());
    CU(configure_subpel());
    ctx->slots.resize(num_slots);
    for (auto& s : ctx->slots) {
        const size_t luma_bytes = align_up((size_t)width * height, 256);
        s.plane_bytes[0] = plane_alloc_bytes(width, height, kPadFull);
        s.plane_bytes[1] = plane_alloc_bytes(width >> 1, height >> 1, kPadQuarter);
        s.plane_bytes[2] = plane_alloc_bytes(width >> 2, height >> 2, kPadSixteenth);
        CU(cudaMalloc(&s.base, luma_bytes + s.plane_bytes[0] + s.plane_bytes[1] + s.plane_bytes[2]));
        s.luma = s.base;
        uint8_t* m = s.base + luma_bytes;
        describe_plane(s.pyr.full, m, width, height, kPadFull); m += s.plane_bytes[0];
        describe_plane(s.pyr.quarter, m, width >> 1, height >> 1, kPadQuarter); m += s.plane_bytes[1];
        describe_plane(s.pyr.sixteenth, m, width >> 2, height >> 2, kPadSixteenth);
        CU(cudaMemsetAsync(s.base, 0, luma_bytes + s.plane_bytes[0] + s.plane_bytes[1] + s.plane_bytes[2], ctx->copy_in));
        CU(cudaEventCreateWithFlags(&s.ready, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&s.ready_ipc, cudaEventDisableTiming | cudaEventInterprocess));
        CU(cudaEventRecord(s.ready, ctx->copy_in));
        CU(cudaEventRecord(s.ready_ipc, ctx->copy_in));
    }
    CU(cudaStreamSynchronize(ctx->copy_in));
    // The full-pel kernels fetch their raw window as ONE 2-D tensor box per SB (TMA) instead of ~1260 16-byte cp.async granules
    // (measured: profiles/microbench/tma_window_r02.txt). One descriptor per slot (the planes never move). SVTME_TMA=0 keeps
    // the cp.async staging; so does a driver without cuTensorMapEncodeTiled unless SVTME_TMA=1 asks for the descriptors.
    const char* tma_env = getenv("SVTME_TMA");
    if (!tma_env || atoi(tma_env) > 0) {
        typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn || q != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            if (tma_env) return fail(ctx, SVTME_ERR_UNDEFINED, "SVTME_TMA=1: cuTensorMapEncodeTiled not available");
            return SVTME_OK;
        }
        std::vector<uint8_t> maps((size_t)num_slots * 128, 0);
        for (int i = 0; i < num_slots; i++) {
            const DPlane& pl = ctx->slots[i].pyr.full;
            const cuuint64_t dims[2] = { (cuuint64_t)pl.stride, (cuuint64_t)(pl.h + 2 * pl.oy) }, strides[1] = { (cuuint64_t)pl.stride };
            const cuuint32_t box[2] = { kTmaBoxBytes, kTmaBoxRows }, estr[2] = { 1, 1 };
            CUtensorMap m;
            memset(&m, 0, sizeof(m));
            const CUresult r = ((EncodeFn)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, pl.buf, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) return fail(ctx, SVTME_ERR_UNDEFINED, "SVTME_TMA: cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
            static_assert(sizeof(CUtensorMap) <= 128, "descriptor slot");
            memcpy(maps.data() + (size_t)i * 128, &m, sizeof(m));
        }
        CU(cudaMalloc(&ctx->d_tmaps, maps.size()));
        CU(cudaMemcpy(ctx->d_tmaps, maps.data(), maps.size(), cudaMemcpyHostToDevice));
    }
    return SVTME_OK;
}

static void free_batch(Batch& b) {
    cudaFree(b.d_scratch); cudaFree(b.d_scratch_ext); cudaFree(b.d_results); cudaFree(b.d_debug); cudaFree(b.d_jobs); cudaFree(b.d_worklist);
    if (b.h_jobs) cudaFreeHost(b.h_jobs);
    b.d_scratch = nullptr; b.d_scratch_ext = nullptr; b.res_pus = 0; b.d_results = nullptr; b.d_debug = nullptr; b.d_jobs = nullptr; b.d_worklist = nullptr; b.h_jobs = nullptr;
    b.cap = 0;
}

void svtme_destroy(SvtMeContext* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    for (cudaStream_t st : { ctx->copy_in, ctx->stream, ctx->copy_out }) if (st) cudaStreamSynchronize(st);
    {   // our peer copies have completed (copy_in is idle): take our events out of the slots of every live peer before they die,
        // and wait for the peers that are still reading OUR slots before the memory goes
        std::lock_guard<std::mutex> live(g_live_mu);
        g_live.erase(ctx);
        for (SvtMeContext* other : g_live) {
            std::lock_guard<std::mutex> lock(other->peer_mu);
            for (auto& s : other->slots)
                for (cudaEvent_t e : ctx->peer_events)
                    s.peer_reads.erase(std::remove(s.peer_reads.begin(), s.peer_reads.end(), e), s.peer_reads.end());
        }
        std::lock_guard<std::mutex> lock(ctx->peer_mu);
        for (auto& s : ctx->slots) {
            for (cudaEvent_t e : s.peer_reads) cudaEventSynchronize(e);
            s.peer_reads.clear();
        }
    }
    for (auto& t : ctx->timed) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
    for (auto e : ctx->event_pool) cudaEventDestroy(e);
    for (auto& s : ctx->slots) { cudaFree(s.base); if (s.ready) cudaEventDestroy(s.ready); if (s.ready_ipc) cudaEventDestroy(s.ready_ipc); }
    cudaFree(ctx->d_tmaps);
    cudaFree(ctx->d_zz);
    cudaFree(ctx->d_ois);
    cudaFree(ctx->d_in16); cudaFree(ctx->d_rem);
    for (auto e : ctx->peer_events) cudaEventDestroy(e);
    for (auto& kv : ctx->ipc_open) { cudaIpcCloseMemHandle(kv.second.first); cudaEventDestroy(kv.second.second); }
    for (auto& b : ctx->ring) {
        free_batch(b);
        if (b.computed) cudaEventDestroy(b.computed);
        if (b.drained) cudaEventDestroy(b.drained);
    }
    if (ctx->join_ev) cudaEventDestroy(ctx->join_ev);
    if (ctx->fork_ev) cudaEventDestroy(ctx->fork_ev);
    for (int i = 0; i < SvtMeContext::kChunkStreams; i++) {
        if (ctx->cstream[i]) cudaStreamDestroy(ctx->cstream[i]);
        if (ctx->cjoin_ev[i]) cudaEventDestroy(ctx->cjoin_ev[i]);
    }
    for (cudaStream_t st : { ctx->copy_in, ctx->stream, ctx->copy_out }) if (st) cudaStreamDestroy(st);
    delete ctx;
}

uint32_t svtme_rc_me_distortion(const SvtMePuResult* sb_rows) {      /* EbMotionEstimation.c:7150-7157 */
    uint32_t sum = 0;
    for (int i = 0; i < 16; i++) sum += sb_rows[5 + i].cand[0].distortion;
    return sum;
}

const char* svtme_last_error(const SvtMeContext* ctx) { return ctx ? ctx->err.c_str() : "null context"; }
void* svtme_stream(SvtMeContext* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
int64_t svtme_kernel_launches(const SvtMeContext* ctx) { return ctx ? ctx->launches : 0; }

// A slot may be rewritten only after the last batch that reads it has finished its kernels.
static int32_t guard_slot_rewrite(SvtMeContext* ctx, Slot& s) {
    // The ring entry's event belongs to the slot's last reader only until the entry is reused, kRing submissions later; by then
    // that batch has drained (svtme_motion_estimate_async waits for it before reusing the entry), so no wait is needed at all.
    // (Waiting on the recycled event made the upload of a sequence's NEXT pictures wait for the batches submitted just before
    // it, i.e. uploads did not overlap the searches.)
    if (s.last_reader >= 0 && ctx->seq - s.last_reader_seq <= kRing)
        CU(cudaStreamWaitEvent(ctx->copy_in, ctx->ring[s.last_reader].computed, 0));
    std::lock_guard<std::mutex> lock(ctx->peer_mu);
    for (cudaEvent_t e : s.peer_reads)                    // peers still copying the old picture; a context that has been
        if (cudaStreamWaitEvent(ctx->copy_in, e, 0) != cudaSuccess) (void)cudaGetLastError();   // destroyed since has drained its copies
    s.peer_reads.clear();
    return SVTME_OK;
}

static int32_t build_pyramid(SvtMeContext* ctx, Slot& s, const uint8_t* d_luma, int32_t stride) {
    { LaunchTimer t(ctx, SVTME_KERNEL_PYRAMID, ctx->copy_in); launch_pyramid(d_luma, stride, s.pyr, ctx->copy_in); }
    CU(cudaGetLastError());
    CU(cudaEventRecord(s.ready, ctx->copy_in));
    CU(cudaEventRecord(s.ready_ipc, ctx->copy_in));
    return SVTME_OK;
}

int32_t svtme_picture_upload_device(SvtMeContext* ctx, int32_t slot, const uint8_t* d_luma, int32_t stride) {
    if (!ctx || slot < 0 || slot >= (int)ctx->slots.size() || !d_luma || stride < ctx->width)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_picture_upload_device: bad argument");
    CU(cudaSetDevice(ctx->device));
    Slot& s = ctx->slots[slot];
    int32_t rc = guard_slot_rewrite(ctx, s);
    if (rc != SVTME_OK) return rc;
    return build_pyramid(ctx, s, d_luma, stride);
}

int32_t svtme_pictures_upload_device(SvtMeContext* ctx, int32_t n, const int32_t* slots, const uint8_t* const* d_lumas, int32_t stride) {
    if (!ctx || n <= 0 || n > SVTME_MAX_BATCH || !slots || !d_lumas || stride < ctx->width)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_pictures_upload_device: bad argument");
    for (int i = 0; i < n; i++)
        if (slots[i] < 0 || slots[i] >= (int)ctx->slots.size() || !d_lumas[i])
            return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_pictures_upload_device: bad slot or plane");
    CU(cudaSetDevice(ctx->device));
    PyramidBatch batch;
    for (int i = 0; i < n; i++) {
        Slot& s = ctx->slots[slots[i]];
        int32_t rc = guard_slot_rewrite(ctx, s);
        if (rc != SVTME_OK) return rc;
        batch.luma[i] = d_lumas[i]; batch.base[i] = s.base;
    }
    const Slot& g = ctx->slots[slots[0]];
    { LaunchTimer t(ctx, SVTME_KERNEL_PYRAMID, ctx->copy_in); launch_pyramid_batch(batch, n, stride, g.pyr, g.base, ctx->copy_in); }
    CU(cudaGetLastError());
    for (int i = 0; i < n; i++) {
        Slot& s = ctx->slots[slots[i]];
        CU(cudaEventRecord(s.ready, ctx->copy_in));
        CU(cudaEventRecord(s.ready_ipc, ctx->copy_in));
    }
    return SVTME_OK;
}

int32_t svtme_pictures_upload(SvtMeContext* ctx, int32_t n, const int32_t* slots, const uint8_t* const* lumas, int32_t stride) {
    if (!ctx || n <= 0 || n > SVTME_MAX_BATCH || !slots || !lumas || stride < ctx->width)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_pictures_upload: bad argument");
    for (int i = 0; i < n; i++)
        if (slots[i] < 0 || slots[i] >= (int)ctx->slots.size() || !lumas[i])
            return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_pictures_upload: bad slot or plane");
    CU(cudaSetDevice(ctx->device));
    PyramidBatch batch;
    for (int i = 0; i < n; i++) {
        Slot& s = ctx->slots[slots[i]];
        int32_t rc = guard_slot_rewrite(ctx, s);
        if (rc != SVTME_OK) return rc;
        if (stride == ctx->width)
            CU(cudaMemcpyAsync(s.luma, lumas[i], (size_t)ctx->width * ctx->height, cudaMemcpyHostToDevice, ctx->copy_in));
        else
            CU(cudaMemcpy2DAsync(s.luma, ctx->width, lumas[i], stride, ctx->width, ctx->height, cudaMemcpyHostToDevice, ctx->copy_in));
        batch.luma[i] = s.luma; batch.base[i] = s.base;
    }
    const Slot& g = ctx->slots[slots[0]];
    { LaunchTimer t(ctx, SVTME_KERNEL_PYRAMID, ctx->copy_in); launch_pyramid_batch(batch, n, ctx->width, g.pyr, g.base, ctx->copy_in); }
    CU(cudaGetLastError());
    for (int i = 0; i < n; i++) {
        Slot& s = ctx->slots[slots[i]];
        CU(cudaEventRecord(s.ready, ctx->copy_in));
        CU(cudaEventRecord(s.ready_ipc, ctx->copy_in));
    }
    return SVTME_OK;
}

int32_t svtme_picture_upload(SvtMeContext* ctx, int32_t slot, const uint8_t* luma, int32_t stride) {
    if (!ctx || slot < 0 || slot >= (int)ctx->slots.size() || !luma || stride < ctx->width)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_picture_upload: bad argument");
    CU(cudaSetDevice(ctx->device));
    Slot& s = ctx->slots[slot];
    int32_t rc = guard_slot_rewrite(ctx, s);
    if (rc != SVTME_OK) return rc;
    if (stride == ctx->width)                             // contiguous plane: one linear copy
        CU(cudaMemcpyAsync(s.luma, luma, (size_t)ctx->width * ctx->height, cudaMemcpyHostToDevice, ctx->copy_in));
    else
        CU(cudaMemcpy2DAsync(s.luma, ctx->width, luma, stride, ctx->width, ctx->height, cudaMemcpyHostToDevice, ctx->copy_in));
    return build_pyramid(ctx, s, s.luma, ctx->width);
}

int32_t svtme_picture_upload_16bit(SvtMeContext* ctx, int32_t slot, const uint16_t* luma16, int32_t stride, uint8_t* bit_inc) {
    if (!ctx || slot < 0 || slot >= (int)ctx->slots.size() || !luma16 || stride < ctx->width)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_picture_upload_16bit: bad argument");
    CU(cudaSetDevice(ctx->device));
    Slot& s = ctx->slots[slot];
    int32_t rc = guard_slot_rewrite(ctx, s);
    if (rc != SVTME_OK) return rc;
    const size_t n = (size_t)ctx->width * ctx->height;
    if (!ctx->d_in16) {                                   // one staging plane per context; the upload stream serialises its use
        CU(cudaMalloc(&ctx->d_in16, n * sizeof(uint16_t)));
        CU(cudaMalloc(&ctx->d_rem, n));
    }
    if (stride == ctx->width)
        CU(cudaMemcpyAsync(ctx->d_in16, luma16, n * sizeof(uint16_t), cudaMemcpyHostToDevice, ctx->copy_in));
    else
        CU(cudaMemcpy2DAsync(ctx->d_in16, (size_t)ctx->width * 2, luma16, (size_t)stride * 2, (size_t)ctx->width * 2, ctx->height,
                             cudaMemcpyHostToDevice, ctx->copy_in));
    launch_unpack16(ctx->d_in16, ctx->width, s.luma, bit_inc ? ctx->d_rem : nullptr, ctx->width, ctx->height, ctx->copy_in);
    ctx->launches++;
    CU(cudaGetLastError());
    if (bit_inc) CU(cudaMemcpyAsync(bit_inc, ctx->d_rem, n, cudaMemcpyDeviceToHost, ctx->copy_in));
    return build_pyramid(ctx, s, s.luma, ctx->width);
}

// interior of the padded full plane = the picture itself; every upload flavour leaves it there
static const uint8_t* slot_picture(const Slot& s) { return s.pyr.full.buf + (size_t)s.pyr.full.oy * s.pyr.full.stride + s.pyr.full.ox; }

int32_t svtme_picture_upload_peer(SvtMeContext* ctx, int32_t slot, SvtMeContext* peer, int32_t peer_slot) {
    if (!ctx || !peer || peer == ctx || slot < 0 || slot >= (int)ctx->slots.size() || peer_slot < 0 ||
        peer_slot >= (int)peer->slots.size() || peer->width != ctx->width || peer->height != ctx->height)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_picture_upload_peer: bad argument");
    CU(cudaSetDevice(ctx->device));
    bool direct = true;                                   // peer memory is addressable from our kernels
    if (peer->device != ctx->device) {
        int can = 0;
        CU(cudaDeviceCanAccessPeer(&can, ctx->device, peer->device));
        if (can) {
            const cudaError_t e = cudaDeviceEnablePeerAccess(peer->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CU(e);
            (void)cudaGetLastError();                     // clear "already enabled"
        }
        direct = can != 0;
    }
    Slot& s = ctx->slots[slot];
    Slot& ps = peer->slots[peer_slot];
    int32_t rc = guard_slot_rewrite(ctx, s);
    if (rc != SVTME_OK) return rc;
    CU(cudaStreamWaitEvent(ctx->copy_in, ps.ready, 0));   // the peer's upload (and pyramid) of that picture
    if (direct) {
        // the pyramid kernel reads the peer's picture straight over NVLink: no staging copy, ~1.3 reads per pixel
        rc = build_pyramid(ctx, s, slot_picture(ps), ps.pyr.full.stride);
    } else {
        cudaMemcpy3DPeerParms cp;                         // no peer access: the driver stages the copy
        memset(&cp, 0, sizeof(cp));
        cp.srcDevice = peer->device; cp.dstDevice = ctx->device;
        cp.srcPtr = make_cudaPitchedPtr(const_cast<uint8_t*>(slot_picture(ps)), ps.pyr.full.stride, ctx->width, ctx->height);
        cp.dstPtr = make_cudaPitchedPtr(s.luma, ctx->width, ctx->width, ctx->height);
        cp.extent = make_cudaExtent(ctx->width, ctx->height, 1);
        CU(cudaMemcpy3DPeerAsync(&cp, ctx->copy_in));
        rc = build_pyramid(ctx, s, s.luma, ctx->width);
    }
    if (rc != SVTME_OK) return rc;
    if (ctx->peer_events.size() < 64) {
        cudaEvent_t e = nullptr;
        CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ctx->peer_events.push_back(e);
    }
    cudaEvent_t done = ctx->peer_events[ctx->next_peer_event++ % ctx->peer_events.size()];
    CU(cudaEventRecord(done, ctx->copy_in));              // the peer's next rewrite of that slot waits for this
    { std::lock_guard<std::mutex> lock(peer->peer_mu); ps.peer_reads.push_back(done); }
    return SVTME_OK;
}

int32_t svtme_slot_export(SvtMeContext* ctx, int32_t slot, SvtMeIpcSlot* out) {
    if (!ctx || !out || slot < 0 || slot >= (int)ctx->slots.size()) return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_slot_export: bad argument");
    CU(cudaSetDevice(ctx->device));
    const Slot& s = ctx->slots[slot];
    static_assert(sizeof(cudaIpcMemHandle_t) == 64 && sizeof(cudaIpcEventHandle_t) == 64, "token layout");
    memset(out, 0, sizeof(*out));
    CU(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->mem_handle), s.base));
    CU(cudaIpcGetEventHandle(reinterpret_cast<cudaIpcEventHandle_t*>(out->event_handle), s.ready_ipc));
    out->luma_offset = (int64_t)(slot_picture(s) - s.base);
    out->pitch = s.pyr.full.stride; out->device = ctx->device; out->width = ctx->width; out->height = ctx->height;
    return SVTME_OK;
}

int32_t svtme_synchronize_uploads(SvtMeContext* ctx) {
    if (!ctx) return SVTME_ERR_BAD_PARAMETER;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->copy_in));
    return SVTME_OK;
}

int32_t svtme_picture_upload_ipc(SvtMeContext* ctx, int32_t slot, const SvtMeIpcSlot* src) {
    if (!ctx || !src || slot < 0 || slot >= (int)ctx->slots.size() || src->width != ctx->width || src->height != ctx->height)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_picture_upload_ipc: bad argument");
    {   // the token must describe a slot of a context with this geometry: same pitch and luma offset as our own slots
        const Slot& own = ctx->slots[slot];
        if (src->pitch != own.pyr.full.stride || src->luma_offset != (int64_t)(slot_picture(own) - own.base))
            return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_picture_upload_ipc: token pitch / offset do not match this geometry");
    }
    CU(cudaSetDevice(ctx->device));
    const std::string key(reinterpret_cast<const char*>(src->mem_handle), 64);
    auto it = ctx->ipc_open.find(key);
    if (it == ctx->ipc_open.end()) {
        void* base = nullptr; cudaEvent_t ev = nullptr;
        cudaIpcMemHandle_t mh; cudaIpcEventHandle_t eh;
        memcpy(&mh, src->mem_handle, 64); memcpy(&eh, src->event_handle, 64);
        CU(cudaIpcOpenMemHandle(&base, mh, cudaIpcMemLazyEnablePeerAccess));
        CU(cudaIpcOpenEventHandle(&ev, eh));
        it = ctx->ipc_open.emplace(key, std::make_pair(base, ev)).first;
    }
    Slot& s = ctx->slots[slot];
    int32_t rc = guard_slot_rewrite(ctx, s);
    if (rc != SVTME_OK) return rc;
    CU(cudaStreamWaitEvent(ctx->copy_in, it->second.second, 0));
    // the mapping is peer memory our kernels can address: the pyramid kernel reads the picture straight over NVLink
    return build_pyramid(ctx, s, static_cast<const uint8_t*>(it->second.first) + src->luma_offset, src->pitch);
}

int32_t svtme_picture_download(SvtMeContext* ctx, int32_t slot, int32_t level, uint8_t* dst, size_t dst_bytes,
                               int32_t* out_stride, int32_t* out_rows, int32_t* out_pad) {
    if (!ctx || slot < 0 || slot >= (int)ctx->slots.size() || level < 0 || level > 2 || !dst)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_picture_download: bad argument");
    const DPlane& pl = level == 0 ? ctx->slots[slot].pyr.full : level == 1 ? ctx->slots[slot].pyr.quarter : ctx->slots[slot].pyr.sixteenth;
    const int w = pl.w + 2 * pl.ox, rows = pl.h + 2 * pl.oy;
    if (dst_bytes < (size_t)w * rows) return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_picture_download: buffer too small");
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpy2DAsync(dst, w, pl.buf, pl.stride, w, rows, cudaMemcpyDeviceToHost, ctx->copy_in));
    CU(cudaStreamSynchronize(ctx->copy_in));
    if (out_stride) *out_stride = w;
    if (out_rows) *out_rows = rows;
    if (out_pad) *out_pad = pl.ox;
    return SVTME_OK;
}

// Make ring entry `b` large enough for n pictures. The entry must be idle (its previous batch drained).
static int32_t ensure_batch(SvtMeContext* ctx, Batch& b, int n, bool ext) {
    const int pus = (ext || b.res_pus == kMaxPu) ? kMaxPu : kNumPu;
    if (n <= b.cap && pus == b.res_pus) return SVTME_OK;
    n = n > b.cap ? n : b.cap;
    free_batch(b);
    CU(cudaMalloc(&b.d_scratch, sizeof(SbScratch) * (size_t)n * ctx->n_sb));
    if (pus == kMaxPu) CU(cudaMalloc(&b.d_scratch_ext, sizeof(SbScratchExt) * (size_t)n * ctx->n_sb));
    CU(cudaMalloc(&b.d_results, sizeof(SvtMePuResult) * (size_t)n * ctx->n_sb * pus));
    b.res_pus = pus;
    CU(cudaMalloc(&b.d_debug, sizeof(SvtMeSbDebug) * (size_t)n * ctx->n_sb));
    CU(cudaMalloc(&b.d_jobs, sizeof(DJob) * (size_t)n));
    CU(cudaMalloc(&b.d_worklist, sizeof(uint32_t) * (size_t)n * (ctx->n_sb + 1)));
    CU(cudaMallocHost(&b.h_jobs, sizeof(DJob) * (size_t)n));
    b.cap = n;
    return SVTME_OK;
}

static bool params_ok(const SvtMeContext* ctx, const SvtMeParams& p, std::string& why) {
#define REQ(c) if (!(c)) { why = "unsupported SvtMeParams: " #c; return false; }
    REQ(p.width == ctx->width && p.height == ctx->height);
    REQ(p.num_lists == 1 || p.num_lists == 2);
    REQ(p.num_pus == SVTME_SQUARE_PU_COUNT || p.num_pus == SVTME_MAX_PU_COUNT);
    REQ(p.hierarchical_levels >= 0 && p.hierarchical_levels <= 5);
    REQ(p.temporal_layer_index >= 0 && p.temporal_layer_index <= p.hierarchical_levels);
    REQ(p.num_hme_regions_w >= 1 && p.num_hme_regions_w <= 2 && p.num_hme_regions_h >= 1 && p.num_hme_regions_h <= 2);
    /* the reference keeps the search area in a uint8_t (EbMotionEstimationContext.h:403-404) and searches MIN(area, 127) */
    REQ(p.search_area_w >= 1 && p.search_area_w <= 255 && p.search_area_h >= 1 && p.search_area_h <= 255);
    REQ(p.hme_l0_mult_x > 0 && p.hme_l0_mult_y > 0);
    for (int i = 0; i < 2; i++) {
        REQ(p.hme_l0_w[i] >= 1 && p.hme_l0_h[i] >= 1 && p.hme_l1_w[i] >= 1 && p.hme_l1_h[i] >= 1 && p.hme_l2_w[i] >= 0 && p.hme_l2_h[i] >= 0);
        REQ(p.hme_l1_w[i] <= 64 && p.hme_l1_h[i] <= 64 && p.hme_l2_w[i] <= 64 && p.hme_l2_h[i] <= 64);
        REQ(p.hme_l0_w[i] * p.hme_l0_mult_x / 100 <= 504 && p.hme_l0_h[i] * p.hme_l0_mult_y / 100 <= 248);   /* argmin key: 9-bit col, 8-bit row */
    }
    REQ(p.use_subpel == 0 || p.use_subpel == 1);
#undef REQ
    return true;
}

int32_t svtme_motion_estimate_async(SvtMeContext* ctx, const SvtMeJob* jobs, int32_t n_jobs) {
    if (!ctx || !jobs || n_jobs <= 0 || n_jobs > SVTME_MAX_BATCH)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_motion_estimate: bad batch");
    CU(cudaSetDevice(ctx->device));
    const int nslots = (int)ctx->slots.size();
    int max_lists = 1;
    bool any_sq = false, any_ext = false;
    bool cls[2][2] = { { false, false }, { false, false } };   // [209-PU mode][search area beyond 64 x 64]: kernel variants this batch needs
    for (int i = 0; i < n_jobs; i++) {
        const SvtMeJob& j = jobs[i];
        (j.params.num_pus == SVTME_MAX_PU_COUNT ? any_ext : any_sq) = true;
        cls[j.params.num_pus == SVTME_MAX_PU_COUNT][wide_search(j.params)] = true;
        std::string why;
        if (!params_ok(ctx, j.params, why)) return fail(ctx, SVTME_ERR_BAD_PARAMETER, why);
        if (j.cur_slot < 0 || j.cur_slot >= nslots || j.ref0_slot < 0 || j.ref0_slot >= nslots ||
            (j.params.num_lists == 2 && (j.ref1_slot < 0 || j.ref1_slot >= nslots)))
            return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_motion_estimate: slot out of range");
        if (j.sb_row_begin != 0 || j.sb_row_end != 0) {
            if (j.sb_row_begin < 0 || j.sb_row_end <= j.sb_row_begin || j.sb_row_end > (j.range_in_sbs ? ctx->n_sb : ctx->sb_rows))
                return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_motion_estimate: SB range out of the picture");
        }
        if (j.range_in_sbs != 0 && j.range_in_sbs != 1) return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_motion_estimate: range_in_sbs must be 0 or 1");
        max_lists = j.params.num_lists > max_lists ? j.params.num_lists : max_lists;
    }
    const int bi = ctx->next;
    Batch& b = ctx->ring[bi];
    if (b.used) CU(cudaEventSynchronize(b.drained));     // the batch that used this entry kRing submissions ago
    int32_t rc = ensure_batch(ctx, b, n_jobs, any_ext);
    if (rc != SVTME_OK) return rc;
    int max_sb = 0;                                       // grid width: the largest SB range of the batch
    size_t res_rows = 0;
    for (int i = 0; i < n_jobs; i++) {
        const SvtMeJob& j = jobs[i];
        DJob& d = b.h_jobs[i];
        const int r1 = j.params.num_lists == 2 ? j.ref1_slot : j.ref0_slot;
        d.cur = ctx->slots[j.cur_slot].pyr;
        d.ref[0] = ctx->slots[j.ref0_slot].pyr;
        d.ref[1] = ctx->slots[r1].pyr;
        d.prm = j.params;
        d.scratch = b.d_scratch + (size_t)i * ctx->n_sb;
        d.scratch_ext = j.params.num_pus == SVTME_MAX_PU_COUNT ? b.d_scratch_ext + (size_t)i * ctx->n_sb : nullptr;
        b.res_off[i] = res_rows;
        d.results = b.d_results + res_rows;
        res_rows += (size_t)ctx->n_sb * j.params.num_pus;
        d.debug = j.debug ? b.d_debug + (size_t)i * ctx->n_sb : nullptr;
        d.sb_cols = ctx->sb_cols; d.sb_rows = ctx->sb_rows;
        d.ref_tmap[0] = ctx->d_tmaps ? (const void*)(ctx->d_tmaps + (size_t)j.ref0_slot * 128) : nullptr;
        d.ref_tmap[1] = ctx->d_tmaps ? (const void*)(ctx->d_tmaps + (size_t)r1 * 128) : nullptr;
        const bool whole = j.sb_row_begin == 0 && j.sb_row_end == 0;
        const int unit = j.range_in_sbs ? 1 : ctx->sb_cols;
        d.sb_begin = whole ? 0 : j.sb_row_begin * unit;
        d.sb_end = whole ? ctx->n_sb : j.sb_row_end * unit;
        max_sb = d.sb_end - d.sb_begin > max_sb ? d.sb_end - d.sb_begin : max_sb;
        for (int sl : { j.cur_slot, j.ref0_slot, r1 }) {
            Slot& s = ctx->slots[sl];
            if (s.last_reader_seq != ctx->seq) {        // first use by this batch: its pyramid must be complete
                CU(cudaStreamWaitEvent(ctx->stream, s.ready, 0));
                s.last_reader = bi; s.last_reader_seq = ctx->seq;
            }
        }
    }
    CU(cudaMemcpyAsync(b.d_jobs, b.h_jobs, sizeof(DJob) * n_jobs, cudaMemcpyHostToDevice, ctx->stream));
    const int n_chunks = ctx->profiling ? 1 : ctx->chunks ? std::min(ctx->chunks, n_jobs) : (n_jobs >= 2 && n_jobs <= 16) ? 2 : 1;
    if (n_chunks > 1) CU(cudaEventRecord(ctx->fork_ev, ctx->stream));
    const int n_cs = std::min(n_chunks, ctx->chunk_streams);
    for (int i = 0; i < (n_chunks > 1 ? n_cs : 0); i++) CU(cudaStreamWaitEvent(ctx->cstream[i], ctx->fork_ev, 0));
    for (int c = 0; c < n_chunks; c++) {
        const int j0 = (int)((int64_t)n_jobs * c / n_chunks), j1 = (int)((int64_t)n_jobs * (c + 1) / n_chunks), nj = j1 - j0;
        const DJob* dj = b.d_jobs + j0;
        cudaStream_t st = n_chunks > 1 ? ctx->cstream[c % n_cs] : ctx->stream;
        for (int list = 0; list < max_lists; list++) {
            { LaunchTimer t(ctx, SVTME_KERNEL_HME, st); launch_hme(dj, nj, max_sb, list, st); }
            // the square (85-PU) / ext (209-PU) and narrow / wide-area variants cover the same grid; a CTA of another class exits at once
            for (int ext = 0; ext < 2; ext++) for (int wide = 0; wide < 2; wide++) {
                if (!cls[ext][wide]) continue;
                { LaunchTimer t(ctx, SVTME_KERNEL_FULLPEL, st); launch_fullpel(dj, nj, max_sb, list, ext != 0, wide != 0, st); }
                { LaunchTimer t(ctx, SVTME_KERNEL_SUBPEL, st); launch_subpel(dj, nj, max_sb, list, ext != 0, wide != 0, b.d_worklist + (size_t)j0 * (ctx->n_sb + 1) + 1, st); }
                if (!wide) ctx->launches++;              // narrow areas: small-window pass + deferred pass
            }
        }
        for (int ext = 0; ext < 2; ext++) {
            if (!(ext ? any_ext : any_sq)) continue;
            LaunchTimer t(ctx, SVTME_KERNEL_BIPRED_PACK, st); launch_bipred_pack(dj, nj, max_sb, ext != 0, st);
        }
    }
    for (int i = 0; i < (n_chunks > 1 ? n_cs : 0); i++) {
        CU(cudaEventRecord(ctx->cjoin_ev[i], ctx->cstream[i]));
        CU(cudaStreamWaitEvent(ctx->stream, ctx->cjoin_ev[i], 0));
    }
    CU(cudaGetLastError());
    CU(cudaEventRecord(b.computed, ctx->stream));
    // results travel on their own stream, straight into the caller's buffers
    CU(cudaStreamWaitEvent(ctx->copy_out, b.computed, 0));
    for (int i = 0; i < n_jobs; i++) {                    // only the SBs of the job's row range travel
        const size_t first = (size_t)i * ctx->n_sb + b.h_jobs[i].sb_begin, count = (size_t)(b.h_jobs[i].sb_end - b.h_jobs[i].sb_begin);
        const size_t pus = (size_t)jobs[i].params.num_pus;
        if (jobs[i].results)
            CU(cudaMemcpyAsync(jobs[i].results, b.d_results + b.res_off[i] + (size_t)b.h_jobs[i].sb_begin * pus, sizeof(SvtMePuResult) * count * pus,
                               cudaMemcpyDeviceToHost, ctx->copy_out));
        if (jobs[i].debug)
            CU(cudaMemcpyAsync(jobs[i].debug, b.d_debug + first, sizeof(SvtMeSbDebug) * count, cudaMemcpyDeviceToHost, ctx->copy_out));
    }
    CU(cudaEventRecord(b.drained, ctx->copy_out));
    b.used = true;
    ctx->last = bi; ctx->seq++;
    ctx->next = (bi + 1) % kRing;
    return SVTME_OK;
}

/* Wait until the results of every batch submitted so far are in the callers' buffers. */
int32_t svtme_fetch_results(SvtMeContext* ctx, const SvtMeJob* jobs, int32_t n_jobs) {
    (void)jobs; (void)n_jobs;
    if (!ctx) return SVTME_ERR_BAD_PARAMETER;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->copy_out));
    return SVTME_OK;
}

int32_t svtme_motion_estimate(SvtMeContext* ctx, const SvtMeJob* jobs, int32_t n_jobs) {
    const int32_t rc = svtme_motion_estimate_async(ctx, jobs, n_jobs);
    if (rc != SVTME_OK) return rc;
    return svtme_fetch_results(ctx, jobs, n_jobs);
}

/* Make the compute stream wait for everything enqueued so far on the transfer streams, so that an event
 * recorded on svtme_stream() afterwards covers uploads, searches and result copies. */
int32_t svtme_join(SvtMeContext* ctx) {
    if (!ctx) return SVTME_ERR_BAD_PARAMETER;
    CU(cudaSetDevice(ctx->device));
    CU(cudaEventRecord(ctx->join_ev, ctx->copy_in));
    CU(cudaStreamWaitEvent(ctx->stream, ctx->join_ev, 0));
    CU(cudaEventRecord(ctx->join_ev, ctx->copy_out));
    CU(cudaStreamWaitEvent(ctx->stream, ctx->join_ev, 0));
    return SVTME_OK;
}

/* ComputeDecimatedZzSad, EbMotionEstimationProcess.c:219-349 (kernel: zz_sad.cu) */
int32_t svtme_zz_sad(SvtMeContext* ctx, const SvtMeZzJob* jobs, int32_t n_jobs) {
    if (!ctx || !jobs || n_jobs <= 0) return ctx ? fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_zz_sad: bad parameter") : SVTME_ERR_BAD_PARAMETER;
    const int n_slots = (int)ctx->slots.size();
    for (int i = 0; i < n_jobs; i++)
        if (jobs[i].cur_slot < 0 || jobs[i].cur_slot >= n_slots || jobs[i].prev_slot < 0 || jobs[i].prev_slot >= n_slots)
            return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_zz_sad: slot out of range");
    CU(cudaSetDevice(ctx->device));
    const size_t jobs_bytes = align_up(sizeof(ZzJob) * (size_t)n_jobs, 256), out_per_job = 2 * (size_t)ctx->n_sb;
    if (n_jobs > ctx->zz_cap) {
        CU(cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->d_zz); ctx->d_zz = nullptr; ctx->zz_cap = 0;
        CU(cudaMalloc(&ctx->d_zz, jobs_bytes + out_per_job * (size_t)n_jobs));
        ctx->zz_cap = n_jobs;
    }
    uint8_t* d_out = ctx->d_zz + align_up(sizeof(ZzJob) * (size_t)ctx->zz_cap, 256);
    std::vector<ZzJob> h((size_t)n_jobs);
    for (int i = 0; i < n_jobs; i++) {
        h[i].cur16 = ctx->slots[jobs[i].cur_slot].pyr.sixteenth;
        h[i].prev16 = ctx->slots[jobs[i].prev_slot].pyr.sixteenth;
        h[i].out = d_out + out_per_job * (size_t)i;
        for (int sl : { jobs[i].cur_slot, jobs[i].prev_slot }) CU(cudaStreamWaitEvent(ctx->stream, ctx->slots[sl].ready, 0));
    }
    CU(cudaMemcpyAsync(ctx->d_zz, h.data(), sizeof(ZzJob) * (size_t)n_jobs, cudaMemcpyHostToDevice, ctx->stream));
    launch_zz_sad(reinterpret_cast<const ZzJob*>(ctx->d_zz), n_jobs, ctx->width, ctx->height, ctx->sb_cols, ctx->n_sb, ctx->stream);
    ctx->launches++;
    CU(cudaGetLastError());
    // one copy for the whole batch (two bytes per SB), scattered into the callers' arrays on the host
    std::vector<uint8_t> host(out_per_job * (size_t)n_jobs);
    CU(cudaMemcpyAsync(host.data(), d_out, host.size(), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));               // synchronous by contract; the slots are free to be rewritten afterwards
    for (int i = 0; i < n_jobs; i++) {
        if (jobs[i].zz_cost) memcpy(jobs[i].zz_cost, host.data() + out_per_job * (size_t)i, (size_t)ctx->n_sb);
        if (jobs[i].non_moving_index) memcpy(jobs[i].non_moving_index, host.data() + out_per_job * (size_t)i + ctx->n_sb, (size_t)ctx->n_sb);
    }
    return SVTME_OK;
}

/* OpenLoopIntraSearchLcu for every SB, EbMotionEstimation.c:8047-8365 (kernel: ois.cu) */
int32_t svtme_ois(SvtMeContext* ctx, const SvtMeOisJob* jobs, int32_t n_jobs) {
    if (!ctx || !jobs || n_jobs <= 0) return ctx ? fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_ois: bad parameter") : SVTME_ERR_BAD_PARAMETER;
    static_assert(sizeof(SvtMeOisParams) == sizeof(ois::Params), "SvtMeOisParams mirrors ois::Params");
    const int n_slots = (int)ctx->slots.size();
    for (int i = 0; i < n_jobs; i++) {
        const SvtMeOisParams& q = jobs[i].params;
        const bool needs_me = q.slice_type != 0 && !(q.temporal_layer_index == 0 && q.input_resolution != 3) && !q.limit_ois_to_dc;
        if (jobs[i].slot < 0 || jobs[i].slot >= n_slots || !jobs[i].cand || !jobs[i].total || q.slice_type < 0 || q.slice_type > 2 ||
            q.temporal_layer_index < 0 || q.temporal_layer_index > 5 || (needs_me && !jobs[i].me_sad))
            return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_ois: bad slot, parameters or missing buffer");
    }
    CU(cudaSetDevice(ctx->device));
    const size_t cus = (size_t)ctx->n_sb * SVTME_OIS_CUS;
    const size_t me_bytes = align_up(cus * 4, 256), cand_bytes = align_up(cus * SVTME_OIS_CANDIDATES * 4, 256), total_bytes = align_up(cus, 256);
    const size_t per_job = me_bytes + cand_bytes + total_bytes;
    if (n_jobs > ctx->ois_cap) {
        CU(cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->d_ois); ctx->d_ois = nullptr; ctx->ois_cap = 0;
        CU(cudaMalloc(&ctx->d_ois, align_up(sizeof(OisJob) * (size_t)n_jobs, 256) + per_job * (size_t)n_jobs));
        ctx->ois_cap = n_jobs;
    }
    uint8_t* base = ctx->d_ois + align_up(sizeof(OisJob) * (size_t)ctx->ois_cap, 256);
    std::vector<OisJob> h((size_t)n_jobs);
    for (int i = 0; i < n_jobs; i++) {
        uint8_t* b = base + per_job * (size_t)i;
        h[i].plane = ctx->slots[jobs[i].slot].pyr.full;
        memcpy(&h[i].prm, &jobs[i].params, sizeof(ois::Params));
        h[i].me_sad = jobs[i].me_sad ? reinterpret_cast<const uint32_t*>(b) : nullptr;
        h[i].out_cand = reinterpret_cast<uint32_t*>(b + me_bytes);
        h[i].out_total = b + me_bytes + cand_bytes;
        CU(cudaStreamWaitEvent(ctx->stream, ctx->slots[jobs[i].slot].ready, 0));
        if (jobs[i].me_sad) CU(cudaMemcpyAsync(b, jobs[i].me_sad, cus * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    CU(cudaMemcpyAsync(ctx->d_ois, h.data(), sizeof(OisJob) * (size_t)n_jobs, cudaMemcpyHostToDevice, ctx->stream));
    { LaunchTimer t(ctx, SVTME_KERNEL_OIS, ctx->stream); launch_ois(reinterpret_cast<const OisJob*>(ctx->d_ois), n_jobs, ctx->sb_cols, ctx->n_sb, ctx->ois_variant, ctx->stream); }
    CU(cudaGetLastError());
    for (int i = 0; i < n_jobs; i++) {
        CU(cudaMemcpyAsync(jobs[i].cand, h[i].out_cand, cus * SVTME_OIS_CANDIDATES * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaMemcpyAsync(jobs[i].total, h[i].out_total, cus, cudaMemcpyDeviceToHost, ctx->stream));
    }
    CU(cudaStreamSynchronize(ctx->stream));               // synchronous by contract
    return SVTME_OK;
}

int32_t svtme_set_profiling(SvtMeContext* ctx, int32_t enable) {
    if (!ctx) return SVTME_ERR_BAD_PARAMETER;
    ctx->profiling = enable != 0;
    return SVTME_OK;
}

int32_t svtme_synchronize(SvtMeContext* ctx) {
    if (!ctx) return SVTME_ERR_BAD_PARAMETER;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->copy_in));
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaStreamSynchronize(ctx->copy_out));
    return SVTME_OK;
}

int32_t svtme_kernel_times(SvtMeContext* ctx, double* ms, int64_t* launches, int32_t reset) {
    if (!ctx) return SVTME_ERR_BAD_PARAMETER;
    const int32_t rc = svtme_synchronize(ctx);
    if (rc != SVTME_OK) return rc;
    for (auto& t : ctx->timed) {
        float e = 0.f;
        CU(cudaEventElapsedTime(&e, t.a, t.b));
        ctx->kind_ms[t.kind] += e;
        ctx->event_pool.push_back(t.a); ctx->event_pool.push_back(t.b);
    }
    ctx->timed.clear();
    for (int k = 0; k < SVTME_KERNEL_KINDS; k++) {
        if (ms) ms[k] = ctx->kind_ms[k];
        if (launches) launches[k] = ctx->kind_launches[k];
        if (reset) { ctx->kind_ms[k] = 0; ctx->kind_launches[k] = 0; }
    }
    return SVTME_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// svtme_motion_estimate_static: the one-picture ME chain with its job table, inter-kernel SbScratch, result rows and
// debug rows in a file-scope device array instead of cudaMalloc'ed ring buffers (for callers that may not allocate:
// CUDA graph capture, tight serving loops). No counterpart in the reference.
//
// g_static_arena is a HOST-SIDE STAND-IN in this file's host code, not a device address: the address the kernels get
// comes from cudaGetSymbolAddress, asked once per device with that device current (static_arena_base). Every context
// of the process on one device shares the array and it keeps the previous call's contents; the call holds
// g_static_mu until its results are on the host, so calls of one process never overlap on the GPU.
struct StaticArena {
    DJob job;
    SbScratch scratch[SVTME_STATIC_MAX_SBS];
    SvtMePuResult results[(size_t)SVTME_STATIC_MAX_SBS * kNumPu];
    SvtMeSbDebug debug[SVTME_STATIC_MAX_SBS];
    uint32_t worklist[SVTME_STATIC_MAX_SBS + 4];      // counter + deferred SBs of the sub-pel passes
};
static __device__ __align__(256) unsigned char g_static_arena[(sizeof(StaticArena) + 255) / 256 * 256];

constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mu;
static StaticArena* g_static_base[kStaticMaxDevices] = {};   // device address of g_static_arena per device

// caller holds g_static_mu and has made ctx->device current
static int32_t static_arena_base(SvtMeContext* ctx, StaticArena** out) {
    if (ctx->device < 0 || ctx->device >= kStaticMaxDevices)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_motion_estimate_static: device ordinal beyond 63");
    if (!g_static_base[ctx->device]) {
        void* p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_static_arena));
        if (!p || ((uintptr_t)p & 255)) return fail(ctx, SVTME_ERR_UNDEFINED, "svtme_motion_estimate_static: bad symbol address");
        g_static_base[ctx->device] = static_cast<StaticArena*>(p);
    }
    *out = g_static_base[ctx->device];
    return SVTME_OK;
}

extern "C" int32_t svtme_motion_estimate_static(SvtMeContext* ctx, const SvtMeJob* job) {
    if (!ctx || !job) return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_motion_estimate_static: null argument");
    const SvtMeJob& j = *job;
    std::string why;
    if (!params_ok(ctx, j.params, why)) return fail(ctx, SVTME_ERR_BAD_PARAMETER, why);
    if (j.params.num_pus != SVTME_SQUARE_PU_COUNT)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_motion_estimate_static: only the 85-PU mode has static scratch (num_pus must be 85)");
    if (ctx->n_sb > SVTME_STATIC_MAX_SBS)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_motion_estimate_static: picture has " + std::to_string(ctx->n_sb) +
                    " SBs, the static scratch holds SVTME_STATIC_MAX_SBS = " + std::to_string(SVTME_STATIC_MAX_SBS));
    const int nslots = (int)ctx->slots.size();
    const int r1 = j.params.num_lists == 2 ? j.ref1_slot : j.ref0_slot;
    if (j.cur_slot < 0 || j.cur_slot >= nslots || j.ref0_slot < 0 || j.ref0_slot >= nslots || r1 < 0 || r1 >= nslots)
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_motion_estimate_static: slot out of range");
    const bool whole = j.sb_row_begin == 0 && j.sb_row_end == 0;
    if (j.range_in_sbs != 0 && j.range_in_sbs != 1) return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_motion_estimate_static: range_in_sbs must be 0 or 1");
    if (!whole && (j.sb_row_begin < 0 || j.sb_row_end <= j.sb_row_begin || j.sb_row_end > (j.range_in_sbs ? ctx->n_sb : ctx->sb_rows)))
        return fail(ctx, SVTME_ERR_BAD_PARAMETER, "svtme_motion_estimate_static: SB range out of the picture");
    CU(cudaSetDevice(ctx->device));
    std::lock_guard<std::mutex> lock(g_static_mu);
    StaticArena* a = nullptr;
    int32_t rc = static_arena_base(ctx, &a);
    if (rc != SVTME_OK) return rc;
    DJob d;
    memset(&d, 0, sizeof(d));
    d.cur = ctx->slots[j.cur_slot].pyr; d.ref[0] = ctx->slots[j.ref0_slot].pyr; d.ref[1] = ctx->slots[r1].pyr;
    d.prm = j.params;
    d.scratch = a->scratch; d.scratch_ext = nullptr; d.results = a->results;
    d.debug = j.debug ? a->debug : nullptr;
    d.sb_cols = ctx->sb_cols; d.sb_rows = ctx->sb_rows;
    d.sb_begin = whole ? 0 : j.sb_row_begin * (j.range_in_sbs ? 1 : ctx->sb_cols);
    d.sb_end = whole ? ctx->n_sb : j.sb_row_end * (j.range_in_sbs ? 1 : ctx->sb_cols);
    const int n_sb = d.sb_end - d.sb_begin;
    cudaStream_t st = ctx->stream;
    for (int sl : { j.cur_slot, j.ref0_slot, r1 }) CU(cudaStreamWaitEvent(st, ctx->slots[sl].ready, 0));
    // the array keeps the previous call's rows: clear the outputs of this call's range so that nothing stale can be
    // mistaken for a result (the kernels themselves write every field they later read; SbScratch needs no clearing)
    CU(cudaMemsetAsync(a->results + (size_t)d.sb_begin * kNumPu, 0, sizeof(SvtMePuResult) * (size_t)n_sb * kNumPu, st));
    if (j.debug) CU(cudaMemsetAsync(a->debug + d.sb_begin, 0, sizeof(SvtMeSbDebug) * (size_t)n_sb, st));
    CU(cudaMemcpyAsync(&a->job, &d, sizeof(DJob), cudaMemcpyHostToDevice, st));
    for (int list = 0; list < j.params.num_lists; list++) {
        launch_hme(&a->job, 1, n_sb, list, st);
        launch_fullpel(&a->job, 1, n_sb, list, false, wide_search(j.params), st);
        launch_subpel(&a->job, 1, n_sb, list, false, wide_search(j.params), a->worklist + 1, st);
        ctx->launches += wide_search(j.params) ? 3 : 4;
    }
    launch_bipred_pack(&a->job, 1, n_sb, false, st);
    ctx->launches++;
    CU(cudaGetLastError());
    if (j.results)
        CU(cudaMemcpyAsync(j.results, a->results + (size_t)d.sb_begin * kNumPu, sizeof(SvtMePuResult) * (size_t)n_sb * kNumPu,
                           cudaMemcpyDeviceToHost, st));
    if (j.debug)
        CU(cudaMemcpyAsync(j.debug, a->debug + d.sb_begin, sizeof(SvtMeSbDebug) * (size_t)n_sb, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));                        // the pictures have been read and the rows are in the caller's buffers
    return SVTME_OK;
}
