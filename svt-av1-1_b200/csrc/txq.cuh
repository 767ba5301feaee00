// txq.cuh -- shared between txq.cu (kernels) and txq_api.cu (C-ABI of include/svttxq.h)
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/svttxq.h"

namespace svtme {

enum { K_DCT = 0, K_ADST = 1, K_FLIPADST = 2, K_IDTX = 3 };      // 1-D class of one pass

struct TxqArgs {
    const int16_t* residual; int stride; int blocks_x; int n_blocks;
    int vk, hk;                              // 1-D class of the vertical / horizontal pass
    SvtTxqQuant q;
    const int16_t* iscan;                    // [pw*ph] inverse scan order of the packed block
    int32_t *coeff, *qcoeff, *dqcoeff; uint16_t* eob; unsigned long long* energy;
};

cudaError_t launch_txq(int txw, int txh, const TxqArgs& a, cudaStream_t st);

}  // namespace svtme
