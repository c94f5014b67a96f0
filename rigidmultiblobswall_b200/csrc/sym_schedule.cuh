// sym_schedule.cuh -- the triangular schedule of the pair-symmetric kernel (uf_sym_kernel), as plain
// host+device functions so that tests/test_sym_schedule.py can enumerate it on a machine with no GPU
// (g++ -DRMB_HOST_ONLY) and check that every unordered pair of 32-blob blocks is visited exactly once.
//
//   row group g  = ROW consecutive blobs (ROW = threads per CTA x blobs per lane), one CTA row;
//                  a launch handles g = row_first + blockIdx.x * row_stride (ranks: row_first = rank, stride = world)
//   tiles        = TILE consecutive records; row g sweeps the tiles [g ROW / TILE, n_tiles), cut into items of
//                  `chunk` tiles (sym_item)
//   blocks       = 32 consecutive blobs; warp w of row g owns blocks I0 .. I0 + NT - 1, I0 = (g warps + w) NT,
//                  and meets block J of a tile as: skip (J < I0) / diagonal / symmetric, per owned block
#pragma once
#include "pair_math.cuh"

namespace rmb {

// One work item = one CTA = one row group x `chunk` consecutive tiles of its triangular row [t_first, n_tiles).
// All items have the same size (the last chunk of a row may be shorter), so the hardware CTA scheduler is the
// dynamic queue and the tail of a launch is at most one item long -- the r01 schedule cut every row into the same
// NUMBER of parts, so the parts of long rows were long (92 ms at N = 1M on one GPU) and the launch drained slowly.
// To keep the grid rectangular, row k of the launch is paired with row rows-1-k (their lengths add up to a
// constant +-1): blockIdx.x = super row k < ceil(rows/2), blockIdx.y = chunk index over the pair; the few grid
// cells past the pair's chunks are invalid and exit at once.
struct SymItem { long g; int v0, v1; bool valid; };   // row group g, tiles [v0, v1)

RMB_HOST_DEVICE int sym_row_len(long g, int row_blobs, int tile_blobs, int n_tiles)
{
  return n_tiles - (int)((g * row_blobs) / tile_blobs);
}

RMB_HOST_DEVICE SymItem sym_item(int bx, int by, int rows, int row_first, int row_stride, int row_blobs, int tile_blobs,
                                 int n_tiles, int chunk)
{
  SymItem it;
  it.valid = false; it.g = 0; it.v0 = it.v1 = 0;
  long g = row_first + (long)bx * row_stride;
  int len = sym_row_len(g, row_blobs, tile_blobs, n_tiles);
  int c = by;
  const int chA = (len + chunk - 1) / chunk;
  if (by >= chA) {
    const int k2 = rows - 1 - bx;
    if (k2 <= bx) return it;                      // the unpaired middle row of an odd launch
    g = row_first + (long)k2 * row_stride;
    len = sym_row_len(g, row_blobs, tile_blobs, n_tiles);
    c = by - chA;
    if (c >= (len + chunk - 1) / chunk) return it;
  }
  it.g = g;
  it.v0 = n_tiles - len + c * chunk;
  it.v1 = it.v0 + chunk < n_tiles ? it.v0 + chunk : n_tiles;
  it.valid = true;
  return it;
}

// blockIdx.y extent: the largest number of chunks of any row pair of the launch (host side, O(rows)).
inline int sym_grid_y(int rows, int row_first, int row_stride, int row_blobs, int tile_blobs, int n_tiles, int chunk)
{
  int y = 0;
  for (int k = 0; 2 * k < rows; k++) {
    const int lenA = sym_row_len(row_first + (long)k * row_stride, row_blobs, tile_blobs, n_tiles);
    int c = (lenA + chunk - 1) / chunk;
    if (rows - 1 - k > k) {
      const int lenB = sym_row_len(row_first + (long)(rows - 1 - k) * row_stride, row_blobs, tile_blobs, n_tiles);
      c += (lenB + chunk - 1) / chunk;
    }
    y = c > y ? c : y;
  }
  return y;
}

// Tiles per item: small enough that every resident CTA slot gets >= `items_per_slot` items (drain of the launch
// <= 1/(2 items_per_slot) of its time), within [chunk_min, chunk_max] (an item must amortise its prologue).
inline int sym_chunk_for(long rows, int row_first, int row_stride, int row_blobs, int tile_blobs, int n_tiles,
                         long resident, int items_per_slot, int chunk_min, int chunk_max)
{
  if (rows <= 0) return chunk_min;
  const double len_first = sym_row_len(row_first, row_blobs, tile_blobs, n_tiles);
  const double len_last = sym_row_len(row_first + (rows - 1) * row_stride, row_blobs, tile_blobs, n_tiles);
  const double tile_rows = 0.5 * (len_first + len_last) * (double)rows;
  long c = (long)(tile_rows / ((double)resident * items_per_slot));
  if (c > chunk_max) c = chunk_max;
  if (c < chunk_min) c = chunk_min;
  return (int)c;
}

// What the owner of block I = I0 + k does with block J: 0 skip, 1 diagonal (one direction, steps 1..31),
// 2 symmetric (32 steps, both directions).
RMB_HOST_DEVICE int sym_block_case(long I, long J) { return J < I ? 0 : (J == I ? 1 : 2); }

}  // namespace rmb
