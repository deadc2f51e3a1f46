// Look-up tables of a block-sparse attention layout, built in one pass on the host.
// layout: int64 [H][M][N] (non-zero = block present).  Outputs, in head-major / row-major / column order (the storage order
// of the compact [batch, nnz, block, block] tensors): head, row-block and column-block index of every non-zero block, and
// a CSR row pointer over the H*M block rows (row_ptr[h*M + m] .. row_ptr[h*M + m + 1] = blocks of that block row), which
// is what the block-sparse softmax and the dsd/dds products walk.
// Reference counterpart: csrc/sparse_attention/utils.cpp (`sdd_segment`, the LUT builder of the Triton SDD matmul).
#include <stdint.h>

#define DSBH_EXPORT extern "C" __attribute__((visibility("default")))

// Returns the number of non-zero blocks.  With out_h == nullptr only counts (and fills row_ptr if given).
DSBH_EXPORT long long dsbh_sparse_lut(const int64_t* layout, int H, int M, int N, int64_t* out_h, int64_t* out_r, int64_t* out_c,
                                      int64_t* row_ptr) {
  long long nnz = 0;
  for (int h = 0; h < H; ++h) {
    for (int m = 0; m < M; ++m) {
      const int64_t* row = layout + (static_cast<long long>(h) * M + m) * N;
      if (row_ptr) row_ptr[static_cast<long long>(h) * M + m] = nnz;
      for (int n = 0; n < N; ++n) {
        if (row[n] != 0) {
          if (out_h) {
            out_h[nnz] = h;
            out_r[nnz] = m;
            out_c[nnz] = n;
          }
          ++nnz;
        }
      }
    }
  }
  if (row_ptr) row_ptr[static_cast<long long>(H) * M] = nnz;
  return nnz;
}
