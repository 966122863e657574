// Host-side helpers that ask the real libstdc++ containers for their iteration order.
//
// Several observable results of the reference depend on std::unordered_map iteration order
// (inventory token order, protocol input/output order, stat reward summation order; see
// SURVEY.md Appendix B5/B6). pybind11's map loader does `clear(); reserve(n); emplace(...)`
// in dict order (pybind11/stl.h map_caster::load), while maps rebuilt inside the core start
// default-constructed. Lowering calls these once per config map and stores lists in the
// resulting order, so the kernels never need a hash table.
#include <cstdint>
#include <cstring>
#include <string>
#include <unordered_map>

extern "C" {

// keys[n] (distinct) inserted in order into unordered_map<uint8_t,int>; out[n] = position in
// `keys` of each element in iteration order. reserve_first mirrors the pybind11 loader.
int mg_umap_order_u8(const uint8_t* keys, int n, int reserve_first, int32_t* out) {
  std::unordered_map<uint8_t, int> m;
  if (reserve_first) m.reserve(static_cast<size_t>(n));
  for (int i = 0; i < n; i++) m.emplace(keys[i], i);
  int k = 0;
  for (const auto& kv : m) out[k++] = kv.second;
  return k;
}

// Same for string keys (stat_rewards: unordered_map<std::string, float>). `keys` is n
// NUL-terminated strings back to back.
int mg_umap_order_str(const char* keys, int n, int reserve_first, int32_t* out) {
  std::unordered_map<std::string, int> m;
  if (reserve_first) m.reserve(static_cast<size_t>(n));
  const char* p = keys;
  for (int i = 0; i < n; i++) {
    m.emplace(std::string(p), i);
    p += std::strlen(p) + 1;
  }
  int k = 0;
  for (const auto& kv : m) out[k++] = kv.second;
  return k;
}

}  // extern "C"
