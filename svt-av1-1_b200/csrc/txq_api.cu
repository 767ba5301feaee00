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
};

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
    TQ(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    TQ(cudaStreamCreateWithFlags(&ctx->copy_in, cudaStreamNonBlocking));
    for (auto& e : ctx->in_done) TQ(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
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
    for (auto& e : ctx->in_done) if (e) cudaEventDestroy(e);
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

int32_t svttxq_synchronize(SvtTxqContext* ctx) {
    if (!ctx) return SVTME_ERR_BAD_PARAMETER;
    TQ(cudaSetDevice(ctx->device));
    TQ(cudaStreamSynchronize(ctx->stream));
    return SVTME_OK;
}

void* svttxq_stream(SvtTxqContext* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
int64_t svttxq_kernel_launches(const SvtTxqContext* ctx) { return ctx ? ctx->launches : 0; }

}  // extern "C"
