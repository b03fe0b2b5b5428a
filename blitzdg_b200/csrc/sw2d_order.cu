// Device element order.  The C ABI takes the mesh in the caller's element order (the reference keeps whatever
// order the .msh file has, src/MeshManager.cpp:130-247); the fused stage kernel works on tiles of 32 consecutive
// elements and pays for every '+' node that lives outside the tile (ghost gathers, phase-1 work, shared memory).
// bdg_sw2d_create therefore renumbers the elements once so that consecutive groups of 32 are compact patches of
// the face-neighbour graph (greedy graph growing, seeds taken from the rim of what is already numbered), uploads
// everything in that order and applies the permutation again in the layout conversions of set_state / get_state /
// rhs -- the caller never sees it, and every element's arithmetic is unchanged (results are bitwise independent of
// the order in per-face LF mode).  On the structured 1000x500 triangle mesh this takes the ghosts per tile from
// 170 (one row of 16 cells) to ~85.
#include <algorithm>
#include <cstdlib>
#include <deque>
#include <vector>

#include "sw2d_common.cuh"

namespace {

// Appends to `order` a permutation (new position -> old element) of the elements [k0,k1); neighbours outside the
// range are ignored.  `first` = size of the first group (so that groups stay aligned with the kernel's tiles when
// k0 is not a multiple of the tile size), then groups of `tile`.
void grow_groups(const int32_t *vmapP, int Np, int Nfp, int Nf, long long K, int k0, int k1, int first, int tile, std::vector<int32_t> &order) {
  const int n = k1 - k0;
  if (n <= 0) return;
  const int Nfn = Nf * Nfp;
  auto nbr = [&](int k, int f) -> int {  // face neighbour inside [k0,k1), else -1
    const long long m = vmapP[(long long)k * Nfn + f * Nfp];
    if (m >= (long long)Np * K) return -1;
    const int kk = (int)(m / Np);
    return (kk == k || kk < k0 || kk >= k1) ? -1 : kk;
  };
  std::vector<uint8_t> state(n, 0);  // 0 free, 1 in the current group's frontier, 2 numbered
  std::vector<int32_t> gain(n, 0);   // faces shared with the current group
  std::deque<int32_t> seeds;         // rim of the numbered region
  std::vector<int32_t> frontier;
  int scan = k0, done = 0;
  int want = first > 0 ? std::min(first, n) : std::min(tile, n);
  while (done < n) {
    int have = 0;
    frontier.clear();
    while (have < want) {
      int pick = -1;
      if (!frontier.empty()) {  // most faces shared with the group; ties: longest in the frontier
        size_t bi = 0;
        for (size_t i = 1; i < frontier.size(); ++i)
          if (gain[frontier[i] - k0] > gain[frontier[bi] - k0]) bi = i;
        pick = frontier[bi];
        frontier.erase(frontier.begin() + bi);
      } else {
        while (!seeds.empty() && state[seeds.front() - k0] == 2) seeds.pop_front();
        if (!seeds.empty()) {
          pick = seeds.front();
          seeds.pop_front();
        } else {
          while (state[scan - k0] == 2) ++scan;
          pick = scan;
        }
      }
      state[pick - k0] = 2;
      order.push_back(pick);
      ++have;
      ++done;
      for (int f = 0; f < Nf; ++f) {
        const int q = nbr(pick, f);
        if (q < 0 || state[q - k0] == 2) continue;
        if (state[q - k0] == 0) {
          state[q - k0] = 1;
          gain[q - k0] = 0;
          frontier.push_back(q);
        }
        ++gain[q - k0];
      }
    }
    std::sort(order.end() - have, order.end());  // inside a group keep the caller's relative order (its regularity helps the bank pattern)
    for (int q : frontier) {  // the group is full: its rim seeds later groups
      state[q - k0] = 0;
      seeds.push_back(q);
    }
    want = std::min(tile, n - done);
  }
}

}  // namespace

// perm[new] = old.  Identity when disabled (BDG_RENUMBER=0) or pointless (K <= tile).  Interior elements
// [0,n_interior) and cut elements [n_interior,K) of a partition are renumbered separately (the split is part of
// the halo contract).
void build_element_order(const int32_t *vmapP, int Np, int Nfp, int Nf, int K, int n_interior, int tile, std::vector<int32_t> &perm) {
  perm.clear();
  const char *env = getenv("BDG_RENUMBER");
  if ((env && env[0] == '0') || K <= tile) return;
  perm.reserve(K);
  if (n_interior < 0 || n_interior > K) n_interior = K;
  grow_groups(vmapP, Np, Nfp, Nf, K, 0, n_interior, 0, tile, perm);
  grow_groups(vmapP, Np, Nfp, Nf, K, n_interior, K, (tile - n_interior % tile) % tile, tile, perm);
}
