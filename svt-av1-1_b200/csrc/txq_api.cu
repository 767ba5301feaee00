// txq_api.cu -- the C-ABI of include/svttxq.h: forward transform + quantisation of batches of transform blocks.
// Host side only; the kernels are in txq.cu. One stream per context; the host-buffer entry point stages through
// device buffers that grow on demand and are kept between calls.
#include "me_common.cuh"
#include "txq.cuh"
#include "../../include/svtme.h"
#include <cstring>
#include <map>
#include <string>
#include <vector>

using namespace svtme;

// TxType -> (vertical, horizontal) 1-D class, in the order of the reference's TxType enum (EbDefinitions.h:
// DCT_DCT, ADST_DCT, DCT_ADST, ADST_ADST, FLIPADST_DCT, DCT_FLIPADST, FLIPADST_FLIPADST, ADST_FLIPADST,
// FLIPADST_ADST, IDTX, V_DCT, H_DCT, V_ADST, H_ADST, V_FLIPADST, H_FLIPADST)
static const uint8_t kVertical[16] = { K_DCT, K_ADST, K_DCT, K_ADST, K_FLIPADST, K_DCT, K_FLIPADST, K_ADST, K_FLIPADST,
                                       K_IDTX, K_DCT, K_IDTX, K_ADST, K_IDTX, K_FLIPADST, K_IDTX };
static const uint8_t kHorizontal[16] = { K_DCT, K_DCT, K_ADST, K_ADST, K_DCT, K_FLIPADST, K_FLIPADST, K_FLIPADST, K_ADST,
                                         K_IDTX, K_IDTX, K_DCT, K_IDTX, K_ADST, K_IDTX, K_FLIPADST };

struct SvtTxqContext {
    int device = 0;
    cudaStream_t stream = nullptr;            // kernels + device-to-host copies
    cudaStream_t copy_in = nullptr;           // host-to-device copies of the host-buffer entry point
    cudaEvent_t in_done[8] = {};              // chunk i of the residual has arrived
    int64_t launches = 0;
    std::string err;
    std::map<int, int16_t*> iscan;         // (txw, txh, scan class) -> device inverse scan table
    // staging for the host-buffer entry point
    void* d_in = nullptr; size_t in_cap = 0;
    void* d_out = nullptr; size_t out_cap = 0;
    // compact return path (svttxq_fwd_txfm_quant_compact): result copies on their own stream, chunk totals in pinned memory
    cudaStream_t copy_out = nullptr;
    cudaEvent_t out_ready[8] = {};            // chunk i: eob / offsets / running total have reached the host
    uint32_t* h_flags = nullptr;              // pinned [8][2]: running total of levels after chunk i, int16 overflow flag
};

namespace svtme {
void launch_txq_compact(const int32_t* d_qcoeff, const uint16_t* d_eob, int n_blocks, int n_coef, const int16_t* d_scan,
                        uint32_t* d_tile_sums, uint32_t* d_offsets, int16_t* d_levels, uint32_t* d_flags, const uint32_t* d_carry, cudaStream_t st);
}

#define TQ(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                         \
            return SVTME_ERR_UNDEFINED;                                                            \
        }                                                                                          \
    } while (0)

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static bool size_ok(int w, int h) {
    auto p2 = [](int v) { return v == 4 || v == 8 || v == 16 || v == 32 || v == 64; };
    if (!p2(w) || !p2(h)) return false;
    const int big = w > h ? w : h, small = w < h ? w : h;
    return big / small <= 4;                                     // the 19 AV1 transform sizes
}

// Scan order of the packed pw x ph block by rule (the reference holds them as tables, av1_scan_orders,
// EbTransforms.h:318-3857): 2-D types walk anti-diagonals -- tall blocks downwards, wide blocks upwards, square
// blocks alternating; V_* types are row-major and H_* types column-major. Returned as the inverse permutation.
static std::vector<int16_t> make_iscan(int txw, int txh, int tx_type) {
    const int pw = txw < 32 ? txw : 32, ph = txh < 32 ? txh : 32;
    std::vector<int16_t> inv((size_t)pw * ph);
    int n = 0;
    if (tx_type >= 10) {
        if ((tx_type & 1) == 0) for (int i = 0; i < pw * ph; i++) inv[i] = (int16_t)n++;
        else for (int c = 0; c < pw; c++) for (int r = 0; r < ph; r++) inv[r * pw + c] = (int16_t)n++;
        return inv;
    }
    for (int d = 0; d < pw + ph - 1; d++) {
        const bool down = ph > pw ? true : ph < pw ? false : (d & 1) != 0;
        const int r0 = d - (pw - 1) > 0 ? d - (pw - 1) : 0, r1 = d < ph - 1 ? d : ph - 1;
        if (down) for (int r = r0; r <= r1; r++) inv[r * pw + (d - r)] = (int16_t)n++;
        else for (int r = r1; r >= r0; r--) inv[r * pw + (d - r)] = (int16_t)n++;
    }
    return inv;
}

static int32_t get_iscan(SvtTxqContext* ctx, int txw, int txh, int tx_type, const int16_t** out) {
    const int cls = tx_type < 10 ? 0 : 1 + (tx_type & 1);
    const int key = (txw << 16) | (txh << 4) | cls;
    auto it = ctx->iscan.find(key);
    if (it == ctx->iscan.end()) {
        const std::vector<int16_t> inv = make_iscan(txw, txh, tx_type);
        std::vector<int16_t> both(inv);                           // inverse scan followed by the scan itself
        both.resize(2 * inv.size());
        for (size_t i = 0; i < inv.size(); i++) both[inv.size() + (size_t)inv[i]] = (int16_t)i;
        int16_t* d = nullptr;
        TQ(cudaMalloc(&d, both.size() * sizeof(int16_t)));
        TQ(cudaMemcpyAsync(d, both.data(), both.size() * sizeof(int16_t), cudaMemcpyHostToDevice, ctx->stream));
        TQ(cudaStreamSynchronize(ctx->stream));                  // `inv` is pageable and dies with this scope
        it = ctx->iscan.emplace(key, d).first;
    }
    *out = it->second;
    return SVTME_OK;
}

extern "C" {

int32_t svttxq_supported(int32_t txw, int32_t txh, int32_t tx_type) {
    if (!size_ok(txw, txh) || tx_type < 0 || tx_type > 15) return 0;
    const int vk = kVertical[tx_type], hk = kHorizontal[tx_type];
    if ((vk == K_ADST || vk == K_FLIPADST) && txh > 16) return 0;
    if ((hk == K_ADST || hk == K_FLIPADST) && txw > 16) return 0;
    if ((txw == 64 || txh == 64) && tx_type != 0) return 0;     // 64-point: DCT_DCT only
    return 1;
}

int32_t svttxq_create(SvtTxqContext** out, int32_t device) {
    if (!out) return SVTME_ERR_BAD_PARAMETER;
    *out = nullptr;
    SvtTxqContext* ctx = new (std::nothrow) SvtTxqContext();
    if (!ctx) return SVTME_ERR_INSUFFICIENT_RESOURCES;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        delete ctx; return SVTME_ERR_UNDEFINED;                 // no GPU: fail loudly, there is no CPU path
    }
    ctx->device = device;
    *out = ctx;
    TQ(cudaSetDevice(device));
    {   // the transform kernels are tiny next to the ME kernels they usually share the GPU with: highest stream priority, so
        // their CTAs are placed as soon as any search CTA retires instead of behind a whole 10 ms grid
        int least = 0, greatest = 0;
        TQ(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        TQ(cudaStreamCreateWithPriority(&ctx->stream, cudaStreamNonBlocking, greatest));
    }
    TQ(cudaStreamCreateWithFlags(&ctx->copy_in, cudaStreamNonBlocking));
    for (auto& e : ctx->in_done) TQ(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    TQ(cudaStreamCreateWithFlags(&ctx->copy_out, cudaStreamNonBlocking));
    for (auto& e : ctx->out_ready) TQ(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    TQ(cudaMallocHost(&ctx->h_flags, 8 * 2 * sizeof(uint32_t)));
    return SVTME_OK;
}

void svttxq_destroy(SvtTxqContext* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (auto& kv : ctx->iscan) cudaFree(kv.second);
    cudaFree(ctx->d_in); cudaFree(ctx->d_out);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    if (ctx->copy_in) cudaStreamDestroy(ctx->copy_in);
    if (ctx->copy_out) { cudaStreamSynchronize(ctx->copy_out); cudaStreamDestroy(ctx->copy_out); }
    for (auto& e : ctx->in_done) if (e) cudaEventDestroy(e);
    for (auto& e : ctx->out_ready) if (e) cudaEventDestroy(e);
    if (ctx->h_flags) cudaFreeHost(ctx->h_flags);
    delete ctx;
}

const char* svttxq_last_error(const SvtTxqContext* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int32_t svttxq_fwd_txfm_quant_device(SvtTxqContext* ctx, const int16_t* d_residual, int32_t stride, int32_t width, int32_t height,
                                     int32_t txw, int32_t txh, int32_t tx_type, const SvtTxqQuant* quant,
                                     int32_t* d_coeff, int32_t* d_qcoeff, int32_t* d_dqcoeff, uint16_t* d_eob, uint64_t* d_energy) {
    if (!ctx) return SVTME_ERR_BAD_PARAMETER;
    if (!d_residual || !quant || width <= 0 || height <= 0 || stride < width || !svttxq_supported(txw, txh, tx_type) ||
        width % txw || height % txh) {
        ctx->err = "svttxq_fwd_txfm_quant: bad parameter"; return SVTME_ERR_BAD_PARAMETER;
    }
    TQ(cudaSetDevice(ctx->device));
    TxqArgs a;
    a.residual = d_residual; a.stride = stride; a.blocks_x = width / txw; a.n_blocks = a.blocks_x * (height / txh);
    a.vk = kVertical[tx_type]; a.hk = kHorizontal[tx_type];
    const int ls = (txw * txh > 256) + (txw * txh > 1024);       // av1_get_tx_scale
    const int rnd = ls ? (1 << (ls - 1)) : 0;
    for (int i = 0; i < 2; i++) {
        TxqLane& l = i ? a.ac : a.dc;
        l.zbin = (int32_t)(int16_t)((uint16_t)(quant->zbin[i] + rnd) >> ls);
        l.round = (int32_t)(int16_t)((uint16_t)(quant->round[i] + rnd) >> ls);
        l.quant_hi = (int32_t)((uint32_t)(int32_t)quant->quant[i] << 16);
        l.qshift = quant->quant_shift[i];
        l.dequant = quant->dequant[i];
    }
    const int32_t rc = get_iscan(ctx, txw, txh, tx_type, &a.iscan);
    if (rc != SVTME_OK) return rc;
    a.scan = a.iscan + (size_t)(txw < 32 ? txw : 32) * (txh < 32 ? txh : 32);
    a.coeff = d_coeff; a.qcoeff = d_qcoeff; a.dqcoeff = d_dqcoeff; a.eob = d_eob; a.energy = (unsigned long long*)d_energy;
    ctx->launches++;
    TQ(launch_txq(txw, txh, a, ctx->stream));
    return SVTME_OK;
}

int32_t svttxq_fwd_txfm_quant(SvtTxqContext* ctx, const int16_t* residual, int32_t stride, int32_t width, int32_t height,
                              int32_t txw, int32_t txh, int32_t tx_type, const SvtTxqQuant* quant,
                              int32_t* coeff, int32_t* qcoeff, int32_t* dqcoeff, uint16_t* eob, uint64_t* energy) {
    if (!ctx) return SVTME_ERR_BAD_PARAMETER;
    if (!residual || !quant || width <= 0 || height <= 0 || stride < width || !svttxq_supported(txw, txh, tx_type) ||
        width % txw || height % txh) {
        ctx->err = "svttxq_fwd_txfm_quant: bad parameter"; return SVTME_ERR_BAD_PARAMETER;
    }
    TQ(cudaSetDevice(ctx->device));
    const size_t n_blocks = (size_t)(width / txw) * (height / txh);
    const size_t per = (size_t)(txw < 32 ? txw : 32) * (txh < 32 ? txh : 32);
    const size_t in_bytes = align_up((size_t)stride * height * sizeof(int16_t), 256);
    const size_t plane = align_up(n_blocks * per * sizeof(int32_t), 256);
    const size_t energy_bytes = align_up(n_blocks * sizeof(uint64_t), 256), eob_bytes = align_up(n_blocks * sizeof(uint16_t), 256);
    const size_t out_bytes = 3 * plane + energy_bytes + eob_bytes;
    if (in_bytes > ctx->in_cap) {
        TQ(cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->d_in); ctx->d_in = nullptr; ctx->in_cap = 0;
        TQ(cudaMalloc(&ctx->d_in, in_bytes)); ctx->in_cap = in_bytes;
    }
    if (out_bytes > ctx->out_cap) {
        TQ(cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->d_out); ctx->d_out = nullptr; ctx->out_cap = 0;
        TQ(cudaMalloc(&ctx->d_out, out_bytes)); ctx->out_cap = out_bytes;
    }
    uint8_t* o = (uint8_t*)ctx->d_out;
    int32_t* d_coeff = (int32_t*)o; int32_t* d_q = (int32_t*)(o + plane); int32_t* d_dq = (int32_t*)(o + 2 * plane);
    uint64_t* d_energy = (uint64_t*)(o + 3 * plane); uint16_t* d_eob = (uint16_t*)(o + 3 * plane + energy_bytes);
    // Block rows go through in up to 8 chunks: the upload of chunk i+1 (copy_in stream) runs against the download of
    // chunk i (main stream), so both directions of the link are busy at once.
    const int block_rows = height / txh, blocks_x = width / txw;
    int n_chunks = (int)(((size_t)width * height) >> 21);             // ~2 Mpixel per chunk
    n_chunks = n_chunks < 1 ? 1 : n_chunks > 8 ? 8 : n_chunks;
    if (n_chunks > block_rows) n_chunks = block_rows;
    for (int c = 0; c < n_chunks; c++) {
        const int r0 = (int)((int64_t)block_rows * c / n_chunks), r1 = (int)((int64_t)block_rows * (c + 1) / n_chunks);
        const size_t first = (size_t)r0 * txh * stride;
        const size_t elems = (size_t)((r1 - r0) * txh - 1) * stride + width;
        TQ(cudaMemcpyAsync((int16_t*)ctx->d_in + first, residual + first, elems * sizeof(int16_t), cudaMemcpyHostToDevice, ctx->copy_in));
        TQ(cudaEventRecord(ctx->in_done[c], ctx->copy_in));
    }
    for (int c = 0; c < n_chunks; c++) {
        const int r0 = (int)((int64_t)block_rows * c / n_chunks), r1 = (int)((int64_t)block_rows * (c + 1) / n_chunks);
        const size_t b0 = (size_t)r0 * blocks_x, nb = (size_t)(r1 - r0) * blocks_x;
        TQ(cudaStreamWaitEvent(ctx->stream, ctx->in_done[c], 0));
        const int32_t rc = svttxq_fwd_txfm_quant_device(ctx, (const int16_t*)ctx->d_in + (size_t)r0 * txh * stride, stride, width, (r1 - r0) * txh,
                                                        txw, txh, tx_type, quant, coeff ? d_coeff + b0 * per : nullptr,
                                                        qcoeff ? d_q + b0 * per : nullptr, dqcoeff ? d_dq + b0 * per : nullptr,
                                                        eob ? d_eob + b0 : nullptr, energy ? d_energy + b0 : nullptr);
        if (rc != SVTME_OK) return rc;
        if (coeff) TQ(cudaMemcpyAsync(coeff + b0 * per, d_coeff + b0 * per, nb * per * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
        if (qcoeff) TQ(cudaMemcpyAsync(qcoeff + b0 * per, d_q + b0 * per, nb * per * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
        if (dqcoeff) TQ(cudaMemcpyAsync(dqcoeff + b0 * per, d_dq + b0 * per, nb * per * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
        if (eob) TQ(cudaMemcpyAsync(eob + b0, d_eob + b0, nb * sizeof(uint16_t), cudaMemcpyDeviceToHost, ctx->stream));
        if (energy) TQ(cudaMemcpyAsync(energy + b0, d_energy + b0, nb * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
    }
    TQ(cudaStreamSynchronize(ctx->stream));
    return SVTME_OK;
}

/* The eob-bounded return path: per block its eob and the eob leading quantised levels in scan order as int16 (what the
 * entropy coder reads; dqcoeff = (level * dequant) >> log_scale is rebuilt by the caller where it needs it,
 * EbFullLoop.c:596-662). Block rows go through in chunks: upload (copy_in), transform + quantisation + prefix sum +
 * gather (stream), then -- once the chunk's running total has reached the host -- the copy of exactly that many levels
 * (copy_out), so both directions of the link stay busy and 2 B per block + 2 B per kept level cross it. */
int32_t svttxq_fwd_txfm_quant_compact(SvtTxqContext* ctx, const int16_t* residual, int32_t stride, int32_t width, int32_t height,
                                      int32_t txw, int32_t txh, int32_t tx_type, const SvtTxqQuant* quant,
                                      uint16_t* eob, uint32_t* level_offset, int16_t* levels, size_t levels_cap, size_t* n_levels) {
    if (!ctx) return SVTME_ERR_BAD_PARAMETER;
    if (!residual || !quant || !eob || !levels || !n_levels || width <= 0 || height <= 0 || stride < width ||
        !svttxq_supported(txw, txh, tx_type) || width % txw || height % txh) {
        ctx->err = "svttxq_fwd_txfm_quant_compact: bad parameter"; return SVTME_ERR_BAD_PARAMETER;
    }
    *n_levels = 0;
    TQ(cudaSetDevice(ctx->device));
    const int block_rows = height / txh, blocks_x = width / txw;
    const size_t n_blocks = (size_t)blocks_x * block_rows;
    const size_t per = (size_t)(txw < 32 ? txw : 32) * (txh < 32 ? txh : 32);
    if (n_blocks * per >= ((size_t)1 << 32)) { ctx->err = "svttxq_fwd_txfm_quant_compact: more than 2^32 coefficients"; return SVTME_ERR_BAD_PARAMETER; }
    const size_t in_bytes = align_up((size_t)stride * height * sizeof(int16_t), 256);
    const size_t q_bytes = align_up(n_blocks * per * sizeof(int32_t), 256), lv_bytes = align_up(n_blocks * per * sizeof(int16_t), 256);
    const size_t eob_bytes = align_up(n_blocks * sizeof(uint16_t), 256), off_bytes = align_up((n_blocks + 9) * sizeof(uint32_t), 256);
    const size_t tile_bytes = align_up((n_blocks / 2048 + 32) * sizeof(uint32_t), 256), flag_bytes = 256;
    const size_t out_bytes = q_bytes + lv_bytes + eob_bytes + off_bytes + tile_bytes + flag_bytes;
    if (in_bytes > ctx->in_cap) {
        TQ(cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->d_in); ctx->d_in = nullptr; ctx->in_cap = 0;
        TQ(cudaMalloc(&ctx->d_in, in_bytes)); ctx->in_cap = in_bytes;
    }
    if (out_bytes > ctx->out_cap) {
        TQ(cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->d_out); ctx->d_out = nullptr; ctx->out_cap = 0;
        TQ(cudaMalloc(&ctx->d_out, out_bytes)); ctx->out_cap = out_bytes;
    }
    uint8_t* o = (uint8_t*)ctx->d_out;
    int32_t* d_q = (int32_t*)o; o += q_bytes;
    int16_t* d_lv = (int16_t*)o; o += lv_bytes;
    uint16_t* d_eob = (uint16_t*)o; o += eob_bytes;
    uint32_t* d_off = (uint32_t*)o; o += off_bytes;
    uint32_t* d_tiles = (uint32_t*)o; o += tile_bytes;
    uint32_t* d_flags = (uint32_t*)o;                                     // [8][2] like h_flags
    const int16_t* d_scan = nullptr;
    {
        const int32_t rc = get_iscan(ctx, txw, txh, tx_type, &d_scan);
        if (rc != SVTME_OK) return rc;
        d_scan += per;                                                    // the scan itself follows its inverse
    }
    int n_chunks = (int)(((size_t)width * height) >> 21);                 // ~2 Mpixel per chunk
    n_chunks = n_chunks < 1 ? 1 : n_chunks > 8 ? 8 : n_chunks;
    if (n_chunks > block_rows) n_chunks = block_rows;
    TQ(cudaMemsetAsync(d_flags, 0, 8 * 2 * sizeof(uint32_t), ctx->stream));
    for (int c = 0; c < n_chunks; c++) {
        const int r0 = (int)((int64_t)block_rows * c / n_chunks), r1 = (int)((int64_t)block_rows * (c + 1) / n_chunks);
        const size_t first = (size_t)r0 * txh * stride;
        const size_t elems = (size_t)((r1 - r0) * txh - 1) * stride + width;
        TQ(cudaMemcpyAsync((int16_t*)ctx->d_in + first, residual + first, elems * sizeof(int16_t), cudaMemcpyHostToDevice, ctx->copy_in));
        TQ(cudaEventRecord(ctx->in_done[c], ctx->copy_in));
    }
    size_t tile0 = 0;
    for (int c = 0; c < n_chunks; c++) {
        const int r0 = (int)((int64_t)block_rows * c / n_chunks), r1 = (int)((int64_t)block_rows * (c + 1) / n_chunks);
        const size_t b0 = (size_t)r0 * blocks_x, nb = (size_t)(r1 - r0) * blocks_x;
        TQ(cudaStreamWaitEvent(ctx->stream, ctx->in_done[c], 0));
        const int32_t rc = svttxq_fwd_txfm_quant_device(ctx, (const int16_t*)ctx->d_in + (size_t)r0 * txh * stride, stride, width, (r1 - r0) * txh,
                                                        txw, txh, tx_type, quant, nullptr, d_q + b0 * per, nullptr, d_eob + b0, nullptr);
        if (rc != SVTME_OK) return rc;
        // offsets continue across chunks: the running total of the chunk before is this chunk's carry (stream order)
        launch_txq_compact(d_q + b0 * per, d_eob + b0, (int)nb, (int)per, d_scan, d_tiles + tile0, d_off + b0 + c, d_lv, d_flags + 2 * c,
                           c ? d_flags + 2 * (c - 1) : nullptr, ctx->stream);
        ctx->launches += 4;
        TQ(cudaGetLastError());
        tile0 += nb / 2048 + 2;
        TQ(cudaMemcpyAsync(eob + b0, d_eob + b0, nb * sizeof(uint16_t), cudaMemcpyDeviceToHost, ctx->stream));
        if (level_offset) TQ(cudaMemcpyAsync(level_offset + b0, d_off + b0 + c, (nb + (c == n_chunks - 1 ? 1 : 0)) * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        TQ(cudaMemcpyAsync(ctx->h_flags + 2 * c, d_flags + 2 * c, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        TQ(cudaEventRecord(ctx->out_ready[c], ctx->stream));
    }
    size_t done = 0;
    bool wide = false;
    for (int c = 0; c < n_chunks; c++) {
        TQ(cudaEventSynchronize(ctx->out_ready[c]));
        const size_t total = ctx->h_flags[2 * c];
        wide |= ctx->h_flags[2 * c + 1] != 0;
        if (total > levels_cap) {
            TQ(cudaStreamSynchronize(ctx->stream)); TQ(cudaStreamSynchronize(ctx->copy_out));
            ctx->err = "svttxq_fwd_txfm_quant_compact: levels buffer too small (" + std::to_string(total) + " levels so far)";
            return SVTME_ERR_INSUFFICIENT_RESOURCES;
        }
        if (total > done) TQ(cudaMemcpyAsync(levels + done, d_lv + done, (total - done) * sizeof(int16_t), cudaMemcpyDeviceToHost, ctx->copy_out));
        done = total;
    }
    TQ(cudaStreamSynchronize(ctx->copy_out));
    TQ(cudaStreamSynchronize(ctx->stream));
    *n_levels = done;
    if (wide) { ctx->err = "svttxq_fwd_txfm_quant_compact: a level does not fit 16 bits (use svttxq_fwd_txfm_quant)"; return SVTME_ERR_BAD_PARAMETER; }
    return SVTME_OK;
}

int32_t svttxq_synchronize(SvtTxqContext* ctx) {
    if (!ctx) return SVTME_ERR_BAD_PARAMETER;
    TQ(cudaSetDevice(ctx->device));
    TQ(cudaStreamSynchronize(ctx->stream));
    return SVTME_OK;
}

void* svttxq_stream(SvtTxqContext* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
int64_t svttxq_kernel_launches(const SvtTxqContext* ctx) { return ctx ? ctx->launches : 0; }

}  // extern "C"
