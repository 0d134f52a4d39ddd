// api_file.inl -- native reader of Lance's legacy (file version 0.1) IVF_PQ `index.idx` (included by api.cu).
// Pure host code: parses the file into the exchange format lb200_index_create takes, so an index written by
// the reference can be opened without any Lance code in the process.  Layout followed:
//   tail  [metadata position u64][major u16][minor u16]"LANC"      rust/lance-io/src/utils.rs:124-156
//   metadata = u32 length + pb::Index                               utils.rs:83-101, rust/lance/src/index.rs:169-181
//   pb::Index / VectorIndex / IVF / PQ / Tensor                     protos/index.proto
//   partition page = n_p*M row-major code bytes, then n_p u64 ids   rust/lance/src/index/vector/pq.rs:292-330

#include <fstream>

struct lb200_index_file {
  int32_t metric = 0;
  uint32_t dim = 0, nlist = 0, M = 0, num_bits = 0, spec_version = 0;
  uint64_t dataset_version = 0;
  std::string name, column;
  std::vector<float> centroids, codebook;
  std::vector<uint64_t> part_offsets, row_ids, page_offsets;
  std::vector<uint32_t> lengths;
  std::vector<uint8_t> codes;
};

namespace {

struct PbReader {
  const uint8_t* p;
  const uint8_t* end;
  bool ok = true;
  PbReader(const uint8_t* b, size_t n) : p(b), end(b + n) {}
  uint64_t varint() {
    uint64_t v = 0;
    for (int shift = 0; shift < 64; shift += 7) {
      if (p >= end) { ok = false; return 0; }
      uint8_t b = *p++;
      v |= (uint64_t)(b & 0x7f) << shift;
      if (!(b & 0x80)) return v;
    }
    ok = false;
    return 0;
  }
  // next field: returns false at the end (or on error: check ok)
  bool next(uint32_t& fno, uint32_t& wt, uint64_t& val, const uint8_t*& data, size_t& len) {
    if (!ok || p >= end) return false;
    uint64_t key = varint();
    if (!ok) return false;
    fno = (uint32_t)(key >> 3);
    wt = (uint32_t)(key & 7);
    data = nullptr;
    len = 0;
    val = 0;
    switch (wt) {
      case 0: val = varint(); break;
      case 1: if (end - p < 8) { ok = false; return false; } memcpy(&val, p, 8); p += 8; break;
      case 5: if (end - p < 4) { ok = false; return false; } { uint32_t t; memcpy(&t, p, 4); val = t; } p += 4; break;
      case 2: {
        uint64_t n = varint();
        if (!ok || (uint64_t)(end - p) < n) { ok = false; return false; }
        data = p;
        len = (size_t)n;
        p += n;
        break;
      }
      default: ok = false; return false;
    }
    return ok;
  }
};

static void pb_packed_u64(uint32_t wt, uint64_t val, const uint8_t* data, size_t len, std::vector<uint64_t>& out, bool& ok) {
  if (wt == 2) {
    PbReader r(data, len);
    while (r.p < r.end && r.ok) out.push_back(r.varint());
    ok = ok && r.ok;
  } else {
    out.push_back(val);
  }
}

static void pb_packed_f32(uint32_t wt, uint64_t val, const uint8_t* data, size_t len, std::vector<float>& out) {
  if (wt == 2) {
    size_t n = len / 4, o = out.size();
    out.resize(o + n);
    memcpy(out.data() + o, data, n * 4);
  } else {
    uint32_t t = (uint32_t)val;
    float f;
    memcpy(&f, &t, 4);
    out.push_back(f);
  }
}

// pb::Tensor -> f32 values + 2-D shape (rust/lance-index/src/vector/utils.rs:125-165)
static int pb_tensor(const uint8_t* b, size_t n, std::vector<float>& out, uint64_t shape[2]) {
  PbReader r(b, n);
  uint32_t fno, wt;
  uint64_t val, dt = 0;
  const uint8_t *data, *payload = nullptr;
  size_t len, plen = 0;
  std::vector<uint64_t> shp;
  bool ok = true;
  while (r.next(fno, wt, val, data, len)) {
    if (fno == 1) dt = val;
    else if (fno == 2) pb_packed_u64(wt, val, data, len, shp, ok);
    else if (fno == 3) { payload = data; plen = len; }
  }
  if (!r.ok || !ok) return set_err(LB200_EINVAL, "malformed tensor in index metadata");
  if (shp.size() != 2) return set_err(LB200_EINVAL, "only accept 2-D tensor shape, got %zu dims", shp.size());
  // shape comes from the file: checked multiplication, and the element count must fit the payload
  if (shp[1] != 0 && shp[0] > (uint64_t)plen / shp[1])
    return set_err(LB200_EINVAL, "Tensor shape [%llu, %llu] does not match to data len: %zu",
                   (unsigned long long)shp[0], (unsigned long long)shp[1], plen);
  const uint64_t cnt = shp[0] * shp[1];
  if (dt == 2) {  // FLOAT32
    if (plen != cnt * 4) return set_err(LB200_EINVAL, "Tensor shape [%llu, %llu] does not match to data len: %zu",
                                        (unsigned long long)shp[0], (unsigned long long)shp[1], plen / 4);
    out.resize(cnt);
    memcpy(out.data(), payload, plen);
  } else if (dt == 3) {  // FLOAT64 -> f32 (the B200 path computes in f32)
    if (plen != cnt * 8) return set_err(LB200_EINVAL, "Tensor shape does not match to data len: %zu", plen / 8);
    out.resize(cnt);
    for (uint64_t i = 0; i < cnt; i++) { double d; memcpy(&d, payload + i * 8, 8); out[i] = (float)d; }
  } else {
    return set_err(LB200_EUNSUPPORTED, "tensor data type %llu is not supported on the B200 path (f32 only)", (unsigned long long)dt);
  }
  shape[0] = shp[0];
  shape[1] = shp[1];
  return 0;
}

static int parse_legacy_index(const uint8_t* blob, size_t size, lb200_index_file* f) {
  if (size < 16) return set_err(LB200_EINVAL, "does not have sufficient data, len: %zu", size);
  if (memcmp(blob + size - 4, "LANC", 4) != 0) return set_err(LB200_EINVAL, "not a Lance file: magic number mismatch");
  uint64_t pos;
  uint16_t major, minor;
  memcpy(&pos, blob + size - 16, 8);
  memcpy(&major, blob + size - 8, 2);
  memcpy(&minor, blob + size - 6, 2);
  if (!(major == 0 && (minor == 1 || minor == 0)))
    return set_err(LB200_EUNSUPPORTED, "index file version %u.%u is not the legacy IVF_PQ layout", major, minor);
  // `pos` comes from the file: compare without additions that can wrap (pos near 2^64)
  if (size < 20 || pos > size - 20) return set_err(LB200_EINVAL, "file size is too small");
  uint32_t mlen;
  memcpy(&mlen, blob + pos, 4);
  if ((uint64_t)mlen > size - 20 - pos) return set_err(LB200_EINVAL, "index metadata runs past the end of the file");

  // ---- pb::Index ----
  PbReader r(blob + pos + 4, mlen);
  uint32_t fno, wt;
  uint64_t val;
  const uint8_t *data, *vec = nullptr;
  size_t len, veclen = 0;
  bool have_col = false;
  while (r.next(fno, wt, val, data, len)) {
    if (fno == 1 && wt == 2) f->name.assign((const char*)data, len);
    else if (fno == 2 && wt == 2 && !have_col) { f->column.assign((const char*)data, len); have_col = true; }
    else if (fno == 3) f->dataset_version = val;
    else if (fno == 5 && wt == 2) { vec = data; veclen = len; }
  }
  if (!r.ok) return set_err(LB200_EINVAL, "malformed index metadata");
  if (!vec) return set_err(LB200_EINVAL, "Index proto was missing implementation field");
  // ---- VectorIndex ----
  PbReader rv(vec, veclen);
  uint64_t metric = 0, pq_dim = 0;
  std::vector<uint64_t> offsets, lengths;
  std::vector<float> cents_rep, cb_rep;
  uint64_t cshape[2] = {0, 0}, bshape[2] = {0, 0};
  bool have_ivf = false, have_pq = false, have_ct = false, have_bt = false, ok = true;
  while (rv.next(fno, wt, val, data, len)) {
    if (fno == 1) f->spec_version = (uint32_t)val;
    else if (fno == 2) f->dim = (uint32_t)val;
    else if (fno == 4) metric = val;
    else if (fno == 3 && wt == 2) {  // VectorIndexStage
      PbReader rs(data, len);
      uint32_t sf, sw;
      uint64_t sv;
      const uint8_t* sd;
      size_t sl;
      while (rs.next(sf, sw, sv, sd, sl)) {
        if (sf == 2 && sw == 2) {  // IVF
          have_ivf = true;
          PbReader ri(sd, sl);
          uint32_t a, b;
          uint64_t c;
          const uint8_t* d;
          size_t l;
          while (ri.next(a, b, c, d, l)) {
            if (a == 1) pb_packed_f32(b, c, d, l, cents_rep);
            else if (a == 2) pb_packed_u64(b, c, d, l, offsets, ok);
            else if (a == 3) pb_packed_u64(b, c, d, l, lengths, ok);
            else if (a == 4 && b == 2) { int rc = pb_tensor(d, l, f->centroids, cshape); if (rc) return rc; have_ct = true; }
          }
          if (!ri.ok) return set_err(LB200_EINVAL, "malformed IVF stage");
        } else if (sf == 3 && sw == 2) {  // PQ
          have_pq = true;
          PbReader rp(sd, sl);
          uint32_t a, b;
          uint64_t c;
          const uint8_t* d;
          size_t l;
          while (rp.next(a, b, c, d, l)) {
            if (a == 1) f->num_bits = (uint32_t)c;
            else if (a == 2) f->M = (uint32_t)c;
            else if (a == 3) pq_dim = c;
            else if (a == 4) pb_packed_f32(b, c, d, l, cb_rep);
            else if (a == 5 && b == 2) { int rc = pb_tensor(d, l, f->codebook, bshape); if (rc) return rc; have_bt = true; }
          }
          if (!rp.ok) return set_err(LB200_EINVAL, "malformed PQ stage");
        } else if (sf == 1) {
          return set_err(LB200_EUNSUPPORTED, "IVF_FLAT sub-index is out of scope of the B200 path");
        } else if (sf == 4) {
          return set_err(LB200_EUNSUPPORTED, "OPQ transform stage is not supported");
        } else if (sf == 5) {
          return set_err(LB200_EUNSUPPORTED, "DiskANN is not supported");
        }
      }
      if (!rs.ok) return set_err(LB200_EINVAL, "malformed index stage");
    }
  }
  if (!rv.ok || !ok) return set_err(LB200_EINVAL, "malformed vector index metadata");
  if (metric > 2) return set_err(LB200_EINVAL, "Metric type '%llu' is not supported", (unsigned long long)metric);
  f->metric = (int32_t)metric;
  if (!have_ivf || !have_pq) return set_err(LB200_EUNSUPPORTED, "index is not IVF_PQ: expected an IVF stage followed by a PQ stage");
  f->nlist = (uint32_t)lengths.size();
  if (!have_ct) {  // deprecated repeated-float centroids (ivf/storage.rs:184-191)
    if (cents_rep.empty()) return set_err(LB200_EINVAL, "IVF stage has no centroids");
    f->centroids.swap(cents_rep);
    cshape[0] = f->nlist;
    cshape[1] = f->nlist ? f->centroids.size() / f->nlist : 0;
  }
  if (!have_bt) f->codebook.swap(cb_rep);
  if (cshape[0] != f->nlist || cshape[1] != f->dim || pq_dim != f->dim)
    return set_err(LB200_EINVAL, "stage dimensions (%llu, %llu) do not match index dimension %u",
                   (unsigned long long)cshape[1], (unsigned long long)pq_dim, f->dim);
  if (!f->M || !f->dim || f->dim % f->M)
    return set_err(LB200_EINVAL, "PQ stage: dimension %u is not divisible by num_sub_vectors %u", f->dim, f->M);
  if (f->num_bits != 8) return set_err(LB200_EUNSUPPORTED, "num_bits %u: only 8-bit PQ is on the B200 path", f->num_bits);
  if (f->codebook.size() != (size_t)256 * f->dim)
    return set_err(LB200_EINVAL, "PQ codebook has %zu values, expected 256x%u", f->codebook.size(), f->dim);
  if (offsets.empty() && !lengths.empty()) {  // ivf/storage.rs:198-207
    uint64_t acc = 0;
    for (uint64_t n : lengths) { offsets.push_back(acc); acc += n; }
  }
  if (offsets.size() != lengths.size()) return set_err(LB200_EINVAL, "IVF stage: offsets and lengths differ in length");

  // ---- partition pages: PQIndex::load of every partition, codes transposed to [M][n_p] ----
  const uint32_t M = f->M;
  f->part_offsets.assign(f->nlist + 1, 0);
  // every size below comes from the file: validate each page against the (validated) metadata position and
  // the running total against the file size BEFORE anything is allocated from them
  const uint64_t max_rows = pos / (M + 8ull);
  for (uint32_t p = 0; p < f->nlist; p++) {
    const uint64_t o = offsets[p], n = lengths[p];
    if (o > pos || n > (pos - o) / (M + 8ull))
      return set_err(LB200_EINVAL, "partition %u page at %llu (%llu rows) overlaps the metadata at %llu", p,
                     (unsigned long long)o, (unsigned long long)n, (unsigned long long)pos);
    if (n > max_rows - f->part_offsets[p])
      return set_err(LB200_EINVAL, "IVF stage: partition lengths add up to more rows than the file can hold");
    f->part_offsets[p + 1] = f->part_offsets[p] + n;
  }
  const uint64_t N = f->part_offsets[f->nlist];
  f->codes.resize((size_t)N * M);
  f->row_ids.resize((size_t)N);
  f->page_offsets = offsets;
  f->lengths.resize(f->nlist);
  for (uint32_t p = 0; p < f->nlist; p++) {
    const uint64_t o = offsets[p], n = lengths[p];
    f->lengths[p] = (uint32_t)n;
    const uint8_t* page = blob + o;
    uint8_t* dst = f->codes.data() + (size_t)f->part_offsets[p] * M;
    for (uint64_t i = 0; i < n; i++)
      for (uint32_t m = 0; m < M; m++) dst[(size_t)m * n + i] = page[i * M + m];  // transpose(), pq/storage.rs:357-377
    if (n) memcpy(f->row_ids.data() + f->part_offsets[p], page + n * M, (size_t)n * 8);
  }
  return 0;
}

}  // namespace

extern "C" int32_t lb200_index_file_open_bytes(const void* bytes, uint64_t size, lb200_index_file** out) {
  if (!out) return set_err(LB200_EINVAL, "null output");
  *out = nullptr;
  if (!bytes && size) return set_err(LB200_EINVAL, "null argument");
  // no C++ exception may cross the C ABI (the host is Rust): allocation failures become error codes
  lb200_index_file* f = nullptr;
  try {
    f = new lb200_index_file();
    int rc = parse_legacy_index((const uint8_t*)bytes, (size_t)size, f);
    if (rc) { delete f; return rc; }
  } catch (const std::bad_alloc&) {
    delete f;
    return set_err(LB200_ENOMEM, "out of host memory while parsing the index file");
  } catch (...) {
    delete f;
    return set_err(LB200_EINVAL, "malformed index file");
  }
  *out = f;
  return 0;
}

extern "C" int32_t lb200_index_file_open(const char* path, lb200_index_file** out) {
  if (!out) return set_err(LB200_EINVAL, "null output");
  *out = nullptr;
  if (!path) return set_err(LB200_EINVAL, "null path");
  try {
    std::ifstream in(path, std::ios::binary | std::ios::ate);
    if (!in) return set_err(LB200_EINVAL, "cannot open index file %s", path);
    const std::streamsize n = in.tellg();
    std::vector<uint8_t> buf((size_t)std::max<std::streamsize>(n, 0));
    in.seekg(0);
    if (n > 0 && !in.read((char*)buf.data(), n)) return set_err(LB200_EINVAL, "cannot read index file %s", path);
    return lb200_index_file_open_bytes(buf.data(), buf.size(), out);
  } catch (const std::bad_alloc&) {
    return set_err(LB200_ENOMEM, "out of host memory while reading %s", path);
  } catch (...) {
    return set_err(LB200_EINVAL, "cannot read index file %s", path);
  }
}

extern "C" int32_t lb200_index_file_get_info(const lb200_index_file* f, lb200_index_file_info* info) {
  if (!f || !info) return set_err(LB200_EINVAL, "null argument");
  info->metric = f->metric;
  info->dim = f->dim;
  info->nlist = f->nlist;
  info->num_sub_vectors = f->M;
  info->num_bits = f->num_bits;
  info->spec_version = f->spec_version;
  info->num_rows = f->part_offsets.empty() ? 0 : f->part_offsets.back();
  info->dataset_version = f->dataset_version;
  info->centroids = f->centroids.data();
  info->codebook = f->codebook.data();
  info->part_offsets = f->part_offsets.data();
  info->codes = f->codes.data();
  info->row_ids = f->row_ids.data();
  info->name = f->name.c_str();
  info->column = f->column.c_str();
  return 0;
}

extern "C" void lb200_index_file_close(lb200_index_file* f) { delete f; }

extern "C" int32_t lb200_index_open(int32_t device_id, const char* path, const float* raw_vectors_or_null,
                                    lb200_index** out) {
  if (!out) return set_err(LB200_EINVAL, "null output");
  *out = nullptr;
  lb200_index_file* f = nullptr;
  int rc = lb200_index_file_open(path, &f);
  if (rc) return rc;
  rc = lb200_index_create(device_id, f->metric, f->dim, f->nlist, f->centroids.data(), f->M, f->num_bits,
                          f->codebook.data(), f->part_offsets.data(), f->codes.data(), f->row_ids.data(),
                          raw_vectors_or_null, out);
  delete f;
  return rc;
}
