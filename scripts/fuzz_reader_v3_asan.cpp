// ASan/UBSan fuzz harness of the native v3 index reader (csrc/api_file_v3.inl + api_file.inl compiled stand-alone):
// truncations, footer edits, metadata / JSON / page-table corruptions of a (index.idx, auxiliary.idx) pair written by
// tests/lance_v3_writer.py.  Run: scripts/fuzz_reader_v3_asan.sh <index.idx> <auxiliary.idx>
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include "../include/lance_b200.h"
static thread_local std::string g_err;
static int set_err(int code, const char* fmt, ...) { char buf[512]; va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap); g_err = buf; return code; }
extern "C" const char* lb200_last_error(void) { return g_err.c_str(); }
extern "C" int32_t lb200_index_create(int32_t, int32_t, uint32_t, uint32_t, const float*, uint32_t, uint32_t, const float*, const uint64_t*, const uint8_t*, const uint64_t*, const float*, lb200_index**) { return set_err(-2, "stub"); }
#include "../lance_b200/csrc/api_file.inl"
#include "../lance_b200/csrc/api_file_v3.inl"
#include <random>
static std::vector<uint8_t> slurp(const char* p) { std::ifstream in(p, std::ios::binary); return std::vector<uint8_t>((std::istreambuf_iterator<char>(in)), {}); }
int main(int argc, char** argv) {
  std::vector<uint8_t> gi = slurp(argv[1]), ga = slurp(argv[2]);
  std::mt19937_64 rng(5); int ok = 0, bad = 0;
  for (int it = 0; it < 30000; it++) {
    std::vector<uint8_t> bi = gi, ba = ga;
    std::vector<uint8_t>& b = (it & 1) ? ba : bi;
    const std::vector<uint8_t>& good = (it & 1) ? ga : gi;
    const int mode = (it / 2) % 6;
    uint64_t cm; memcpy(&cm, good.data() + good.size() - 40, 8);
    uint64_t lo = std::min<uint64_t>(cm, good.size() - 41);
    if (good.size() > 4200) { uint64_t g0; memcpy(&g0, good.data() + good.size() - 40 + 16, 8); uint64_t p0; memcpy(&p0, good.data() + g0, 8); lo = std::min(lo, p0); }
    if (mode == 0) { size_t cut = rng() % good.size(); b.resize(cut); b.insert(b.end(), good.end() - 40, good.end()); }
    else if (mode == 1) { for (int k = 0; k < 1 + (int)(rng() % 4); k++) b[lo + rng() % (good.size() - lo)] = (uint8_t)rng(); }
    else if (mode == 2) { b[good.size() - 40 + rng() % 36] = (uint8_t)rng(); }
    else if (mode == 3) { size_t cut = rng() % good.size(); b.resize(cut); }
    else if (mode == 4) { const uint64_t sz = good.size(); const uint64_t cand[] = {~0ull, ~0ull - 15, sz, sz - 40, sz - 41, 0, 1ull << 63, rng()};
      for (int k = 0; k < 3; k++) if (rng() & 1) { uint64_t v = cand[rng() % 8]; memcpy(b.data() + sz - 40 + 8 * k, &v, 8); }
      if (rng() & 1) { uint32_t v = (uint32_t)rng(); memcpy(b.data() + sz - 16 + 4 * (rng() & 1), &v, 4); } }
    else { size_t at = lo + rng() % (good.size() - 40 - lo), n = 1 + rng() % 9; for (size_t k = 0; k < n && at + k < good.size() - 40; k++) b[at + k] = k + 1 == n ? 0x3f : 0xff; }
    uint8_t* hi = (uint8_t*)malloc(bi.size() ? bi.size() : 1); memcpy(hi, bi.data(), bi.size());
    uint8_t* ha = (uint8_t*)malloc(ba.size() ? ba.size() : 1); memcpy(ha, ba.data(), ba.size());
    lb200_index_file* f = nullptr;
    int rc = lb200_index_file_open_v3_bytes(hi, bi.size(), ha, ba.size(), &f);
    if (rc == 0) { ok++; lb200_index_file_info info; lb200_index_file_get_info(f, &info); lb200_index_file_close(f); } else bad++;
    free(hi); free(ha);
  }
  printf("ok %d rejected %d\n", ok, bad);
  return 0;
}
