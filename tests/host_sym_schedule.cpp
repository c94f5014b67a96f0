// tests/host_sym_schedule.cpp -- TEST ONLY.  Enumerates the product's own schedule functions
// (csrc/sym_schedule.cuh, g++ -DRMB_HOST_ONLY) exactly as uf_sym_kernel walks them and counts how often every
// pair of 32-blob blocks is met: symmetric[I][J] (I < J) and diagonal[I] must all be 1 over the union of the
// ranks' launches.
#include <vector>
#include "../rigidmultiblobswall_b200/csrc/sym_schedule.cuh"
using namespace rmb;

extern "C" int host_sym_schedule(long n_blobs, int world, int nt, int tpb, int tile, long resident, int waves,
                                 int* sym_count /* nb*nb */, int* diag_count /* nb */, int* split_out /* world */)
{
  const int n_tiles = (int)((n_blobs + tile - 1) / tile);
  const long n_pad = (long)n_tiles * tile;
  const long nb = n_pad / 32;
  const int row = tpb * nt, warps = tpb / 32;
  const long rows_all = n_pad / row;
  for (int rank = 0; rank < world; rank++) {
    const long rows = (rows_all - rank + world - 1) / world;
    const int split = sym_split_for(rows, resident, waves, n_tiles);
    split_out[rank] = split;
    for (long bx = 0; bx < rows; bx++) {
      const long g = rank + bx * world;
      int covered = (int)((g * row) / tile);                 // the parts must tile [t_first, n_tiles) without gaps
      for (int y = 0; y < split; y++) {
        const SymRowPart p = sym_row_part(g, row, tile, n_tiles, y, split);
        if (p.v0 != covered || p.v1 < p.v0) return -1;
        covered = p.v1;
        for (int w = 0; w < warps; w++) {
          const long I0 = (g * warps + w) * nt;
          for (int v = p.v0; v < p.v1; v++)
            for (int jb = 0; jb < tile / 32; jb++) {
              const long J = (long)v * (tile / 32) + jb;
              for (int k = 0; k < nt; k++) {
                const int c = sym_block_case(I0 + k, J);
                if (c == 1) diag_count[I0 + k]++;
                if (c == 2) sym_count[(I0 + k) * nb + J]++;
              }
            }
        }
      }
      if (covered != n_tiles) return -2;
    }
  }
  return (int)nb;
}
