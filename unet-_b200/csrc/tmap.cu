// tmap.cu -- see tmap.cuh
#include "tmap.cuh"

namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
}  // namespace

int tmap_init(ub_ctx* ctx) {
  if (g_encode) return UB_OK;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn)
    return ub_fail(ctx, UB_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver (%s)", cudaGetErrorString(e));
  g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  return UB_OK;
}

int tmap_make(ub_ctx* ctx, CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides,
              const uint32_t* box) {
  if (!g_encode) { int r = tmap_init(ctx); if (r) return r; }
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides[i];
#ifdef UB_ACT_BF16
  const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
#else
  const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
#endif
  CUresult r = g_encode(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return ub_fail(ctx, UB_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rank %d)", (int)r, rank);
  return UB_OK;
}

int tmap_pp_rows(ub_ctx* ctx, CUtensorMap* out, const act_t* base, int B, int nchunks, const PPLayout& L,
                 int box_cols, int box_rows, int box_chunks) {
  // the 8 channels of a position and the positions of a row are contiguous: one inner dimension of pitch*8
  // elements, so a box row is a single box_cols*16-byte run (a 16-byte inner box would be one request per position)
  const uint64_t dims[4] = {(uint64_t)L.pitch * 8, (uint64_t)L.H + 2, (uint64_t)nchunks, (uint64_t)B};
  const uint64_t strides[3] = {(uint64_t)L.pitch * 16, (uint64_t)L.Ppl * 16, (uint64_t)nchunks * L.Ppl * 16};
  const uint32_t box[4] = {(uint32_t)box_cols * 8, (uint32_t)box_rows, (uint32_t)box_chunks, 1};
  return tmap_make(ctx, out, base + (size_t)L.G * 8, 4, dims, strides, box);
}

int tmap_pp_flat(ub_ctx* ctx, CUtensorMap* out, const act_t* base, int B, int nchunks, const PPLayout& L,
                 int box_pos, int box_chunks) {
  const uint64_t dims[4] = {8, (uint64_t)L.Ppl, (uint64_t)nchunks, (uint64_t)B};
  const uint64_t strides[3] = {16, (uint64_t)L.Ppl * 16, (uint64_t)nchunks * L.Ppl * 16};
  const uint32_t box[4] = {8, (uint32_t)box_pos, (uint32_t)box_chunks, 1};
  return tmap_make(ctx, out, base, 4, dims, strides, box);
}
