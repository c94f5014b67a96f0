// tests/host_sym_schedule.cpp -- TEST ONLY.  Enumerates the product's own schedule functions
// (csrc/sym_schedule.cuh, g++ -DRMB_HOST_ONLY) exactly as sym_kernel walks them and counts how often every
// pair of 32-blob blocks is met: symmetric[I][J] (I < J) and diagonal[I] must all be 1 over the union of the
// ranks' launches.  Also reports, per rank, the grid, the number of valid items and their tile-rows.
#include <vector>
#include "../rigidmultiblobswall_b200/csrc/sym_schedule.cuh"
using namespace rmb;

extern "C" int host_sym_schedule(long n_blobs, int world, int nt, int tpb, int tile, long resident, int items_per_slot,
                                 int chunk_min, int chunk_max, int* sym_count /* nb*nb */, int* diag_count /* nb */,
                                 int* info /* world * 5: grid x, grid y, chunk, valid items, tile-rows */)
{
  const int n_tiles = (int)((n_blobs + tile - 1) / tile);
  const long n_pad = (long)n_tiles * tile;
  const long nb = n_pad / 32;
  const int row = tpb * nt, warps = tpb / 32;
  const long rows_all = n_pad / row;
  for (int rank = 0; rank < world; rank++) {
    const long rows = (rows_all - rank + world - 1) / world;
    int* o = info + 5 * rank;
    o[0] = o[1] = o[2] = o[3] = o[4] = 0;
    if (rows <= 0) continue;
    const int chunk = sym_chunk_for(rows, rank, world, row, tile, n_tiles, resident, items_per_slot, chunk_min, chunk_max);
    const int gy = sym_grid_y((int)rows, rank, world, row, tile, n_tiles, chunk);
    o[0] = (int)((rows + 1) / 2); o[1] = gy; o[2] = chunk;
    std::vector<int> covered(rows, -1);                       // per row: the next tile expected (items tile the row in order)
    for (int by = 0; by < gy; by++)
      for (int bx = 0; bx < (rows + 1) / 2; bx++) {
        const SymItem it = sym_item(bx, by, (int)rows, rank, world, row, tile, n_tiles, chunk);
        if (!it.valid) continue;
        if ((it.g - rank) % world != 0) return -3;
        const long k = (it.g - rank) / world;
        if (k < 0 || k >= rows) return -4;
        const int t_first = (int)((it.g * row) / tile);
        if (covered[k] < 0) covered[k] = t_first;
        if (it.v0 != covered[k] || it.v1 <= it.v0 || it.v1 - it.v0 > chunk) return -1;
        covered[k] = it.v1;
        o[3]++; o[4] += it.v1 - it.v0;
        if (!sym_count) continue;
        for (int w = 0; w < warps; w++) {
          const long I0 = (it.g * warps + w) * nt;
          for (int v = it.v0; v < it.v1; v++)
            for (int jb = 0; jb < tile / 32; jb++) {
              const long J = (long)v * (tile / 32) + jb;
              for (int kk = 0; kk < nt; kk++) {
                const int c = sym_block_case(I0 + kk, J);
                if (c == 1) diag_count[I0 + kk]++;
                if (c == 2) sym_count[(I0 + kk) * nb + J]++;
              }
            }
        }
      }
    for (long k = 0; k < rows; k++)
      if (covered[k] != n_tiles) return -2;
  }
  return (int)nb;
}
