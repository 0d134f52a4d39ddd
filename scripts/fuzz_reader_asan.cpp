// ASan/UBSan fuzz harness of the native index-file reader (csrc/api_file.inl compiled stand-alone with stubbed error
// plumbing): 20 000 truncations / metadata edits / tail edits of a real index file, each parsed from an exact-size heap
// copy.  Run: scripts/fuzz_reader_asan.sh <index.idx>.  Round 1: clean on two reference-written files (40 000 inputs).
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
#include <random>
int main(int argc, char** argv) {
  std::ifstream in(argv[1], std::ios::binary); std::vector<uint8_t> good((std::istreambuf_iterator<char>(in)), {});
  std::mt19937_64 rng(11); int ok = 0, bad = 0;
  for (int it = 0; it < 20000; it++) {
    std::vector<uint8_t> b = good;
    int mode = it % 6;
    uint64_t pos; memcpy(&pos, good.data() + good.size() - 16, 8);
    if (mode == 0) { size_t cut = rng() % good.size(); b.resize(cut); b.insert(b.end(), good.end() - 16, good.end()); }
    else if (mode == 1) { for (int k = 0; k < 1 + (int)(rng() % 4); k++) b[pos + rng() % (good.size() - pos)] = (uint8_t)rng(); }
    else if (mode == 2) { b[good.size() - 16 + rng() % 12] = (uint8_t)rng(); }
    else if (mode == 3) { size_t cut = rng() % good.size(); b.resize(cut); }
    else if (mode == 4) {  // crafted footer positions (pos + 4 wrapped for pos >= 2^64-4)
      const uint64_t sz = good.size();
      const uint64_t cand[] = {~0ull, ~0ull - 3, ~0ull - 19, sz - 16, sz - 19, sz - 20, sz, 1ull << 63, rng()};
      uint64_t p2 = cand[rng() % 9]; memcpy(b.data() + sz - 16, &p2, 8); }
    else {  // saturate a run of metadata bytes: varints (lengths, offsets, tensor shapes) become huge
      size_t at = pos + rng() % (good.size() - 16 - pos), n = 1 + rng() % 9;
      for (size_t k = 0; k < n && at + k < good.size() - 16; k++) b[at + k] = k + 1 == n ? 0x3f : 0xff; }
    // exact-size heap copy so ASan sees any overread
    uint8_t* heap = (uint8_t*)malloc(b.size() ? b.size() : 1); memcpy(heap, b.data(), b.size());
    lb200_index_file* f = nullptr;
    int rc = lb200_index_file_open_bytes(heap, b.size(), &f);
    if (rc == 0) { ok++; lb200_index_file_info info; lb200_index_file_get_info(f, &info); lb200_index_file_close(f); } else bad++;
    free(heap);
  }
  printf("ok %d rejected %d\n", ok, bad);
  return 0;
}
