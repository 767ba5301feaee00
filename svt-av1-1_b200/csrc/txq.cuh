// txq.cuh -- shared between txq.cu (kernels) and txq_api.cu (C-ABI of include/svttxq.h)
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/svttxq.h"

namespace svtme {

enum { K_DCT = 0, K_ADST = 1, K_FLIPADST = 2, K_IDTX = 3 };      // 1-D class of one pass

// One lane (DC or AC) of the quantiser with the transform-size scaling already applied on the host
// (aom_highbd_quantize_b_32x32/_64x64: zbin and round are rounded down by log_scale first).
struct TxqLane { int32_t zbin, round, quant_hi, qshift, dequant; };     // quant_hi = quant << 16: (t * quant) >> 16 == mulhi(t, quant_hi)

struct TxqArgs {
    const int16_t* residual; int stride; int blocks_x; int n_blocks;
    int vk, hk;                              // 1-D class of the vertical / horizontal pass
    TxqLane dc, ac;
    const int16_t* iscan;                    // [pw*ph] inverse scan order of the packed block
    const int16_t* scan;                     // [pw*ph] scan order (position -> raster index)
    int32_t *coeff, *qcoeff, *dqcoeff; uint16_t* eob; unsigned long long* energy;
};

cudaError_t launch_txq(int txw, int txh, const TxqArgs& a, cudaStream_t st);

}  // namespace svtme
