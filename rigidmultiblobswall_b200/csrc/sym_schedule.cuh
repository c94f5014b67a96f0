// sym_schedule.cuh -- the triangular schedule of the pair-symmetric kernel (uf_sym_kernel), as plain
// host+device functions so that tests/test_sym_schedule.py can enumerate it on a machine with no GPU
// (g++ -DRMB_HOST_ONLY) and check that every unordered pair of 32-blob blocks is visited exactly once.
//
//   row group g  = ROW consecutive blobs (ROW = threads per CTA x blobs per lane), one CTA row;
//                  a launch handles g = row_first + blockIdx.x * row_stride (ranks: row_first = rank, stride = world)
//   tiles        = TILE consecutive records; row g sweeps the tiles [g ROW / TILE, n_tiles), cut into n_split
//                  contiguous parts (blockIdx.y)
//   blocks       = 32 consecutive blobs; warp w of row g owns blocks I0 .. I0 + NT - 1, I0 = (g warps + w) NT,
//                  and meets block J of a tile as: skip (J < I0) / diagonal / symmetric, per owned block
#pragma once
#include "pair_math.cuh"

namespace rmb {

struct SymRowPart { int v0, v1; };   // tiles [v0, v1)

RMB_HOST_DEVICE SymRowPart sym_row_part(long g, int row_blobs, int tile_blobs, int n_tiles, int y, int n_split)
{
  const int t_first = (int)((g * row_blobs) / tile_blobs);
  const int len = n_tiles - t_first;
  SymRowPart p;
  p.v0 = t_first + (int)(((long)len * y) / n_split);
  p.v1 = t_first + (int)(((long)len * (y + 1)) / n_split);
  return p;
}

// blockIdx.y extent of one launch: rows are triangular (1 .. n_tiles tiles), so a launch (one rank's share of the
// rows) is split until it has `waves` waves of the `resident` CTAs the chip holds; a part keeps >= 4 tiles.
RMB_HOST_DEVICE int sym_split_for(long rows, long resident, int waves, int n_tiles)
{
  long split = rows > 0 ? (waves * resident + rows - 1) / rows : 1;
  if (split > n_tiles / 4) split = n_tiles / 4;
  return split < 1 ? 1 : (int)split;
}

// What the owner of block I = I0 + k does with block J: 0 skip, 1 diagonal (one direction, steps 1..31),
// 2 symmetric (32 steps, both directions).
RMB_HOST_DEVICE int sym_block_case(long I, long J) { return J < I ? 0 : (J == I ? 1 : 2); }

}  // namespace rmb
