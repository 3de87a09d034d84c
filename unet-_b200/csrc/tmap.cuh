// tmap.cuh -- TMA tensor maps (cuTensorMapEncodeTiled, resolved at run time so the library
// links against libcudart only) over the pixel-planar activation layout of common.cuh.
#pragma once
#include <cuda.h>
#include "common.cuh"

// Resolves the driver entry point once per process; error text goes to ub_last_error.
int tmap_init(ub_ctx* ctx);
// Generic tiled map of 16-bit activations: rank <= 5, dims/box innermost first, strides in bytes
// for dims 1..rank-1.
int tmap_make(ub_ctx* ctx, CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
              const uint64_t* strides, const uint32_t* box);
// [b][chunk][padded row][padded col][8 ch] view of a PP tensor; box = box_rows x box_cols positions of
// box_chunks chunks (box_cols <= 32).  Coordinates: (8*col, row, chunk, b) in padded image coordinates.
int tmap_pp_rows(ub_ctx* ctx, CUtensorMap* out, const act_t* base, int B, int nchunks, const PPLayout& L,
                 int box_cols, int box_rows, int box_chunks);
// [b][chunk][position][8 ch] view; box = box_pos consecutive positions of box_chunks chunks.
// Coordinates: (0, position, chunk, b).
int tmap_pp_flat(ub_ctx* ctx, CUtensorMap* out, const act_t* base, int B, int nchunks, const PPLayout& L,
                 int box_pos, int box_chunks);
