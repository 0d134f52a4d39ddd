// api_file_v3.inl -- native reader of Lance's CURRENT IVF_PQ index (IndexFileVersion::V3: index.idx + auxiliary.idx,
// both Lance-v2 file containers; included by api.cu).  Pure host code.
// PARITY UNPINNED: the reference tree holds no v3 index file and its Rust writer cannot run in this image; this parser
// and the Python one (lance_b200/reader_v3.py, which documents the layout source by source) are pinned on each other
// and on files written by tests/lance_v3_writer.py.  Layout followed:
//   container   protos/file2.proto, rust/lance-file/src/v2/writer.rs:395-590 (footer, CMO / GBO tables, schema in
//               global buffer 0 as pb::FileDescriptor)
//   encodings   protos/encodings.proto: Nullable{no_nulls{Flat}} and FixedSizeList{dim, Nullable{no_nulls{Flat 8}}}
//   contents    rust/lance/src/index/vector/builder.rs:624-770 (writer), rust/lance/src/index/vector/ivf/v2.rs:110-202,
//               rust/lance-index/src/vector/storage.rs:210-298 (readers), rust/lance-index/src/vector/pq.rs:438-457
//               (ProductQuantizationMetadata JSON)

#include <map>

namespace {

// ---- a small JSON reader: what serde_json writes for IndexMetadata / ProductQuantizationMetadata ----
struct Json {
  enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
  bool b = false;
  double num = 0;
  std::string str;
  std::vector<Json> arr;
  std::vector<std::pair<std::string, Json>> obj;
  const Json* get(const char* k) const {
    for (auto& kv : obj)
      if (kv.first == k) return &kv.second;
    return nullptr;
  }
};

struct JsonParser {
  const char* p;
  const char* end;
  int depth = 0;
  bool ok = true;
  void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) p++; }
  bool lit(const char* s) {
    size_t n = strlen(s);
    if ((size_t)(end - p) < n || memcmp(p, s, n) != 0) return false;
    p += n;
    return true;
  }
  Json value() {
    Json j;
    if (++depth > 64) { ok = false; return j; }
    ws();
    if (p >= end) { ok = false; depth--; return j; }
    if (*p == '{') {
      j.kind = Json::Obj;
      p++;
      ws();
      if (p < end && *p == '}') { p++; depth--; return j; }
      while (ok) {
        ws();
        if (p >= end || *p != '"') { ok = false; break; }
        Json k = value();
        ws();
        if (!ok || p >= end || *p != ':') { ok = false; break; }
        p++;
        Json v = value();
        j.obj.emplace_back(std::move(k.str), std::move(v));
        ws();
        if (p < end && *p == ',') { p++; continue; }
        if (p < end && *p == '}') { p++; break; }
        ok = false;
      }
    } else if (*p == '[') {
      j.kind = Json::Arr;
      p++;
      ws();
      if (p < end && *p == ']') { p++; depth--; return j; }
      while (ok) {
        j.arr.push_back(value());
        ws();
        if (p < end && *p == ',') { p++; continue; }
        if (p < end && *p == ']') { p++; break; }
        ok = false;
      }
    } else if (*p == '"') {
      j.kind = Json::Str;
      p++;
      while (p < end && *p != '"') {
        if (*p == '\\') {
          if (end - p < 2) { ok = false; break; }
          char c = p[1];
          p += 2;
          switch (c) {
            case 'n': j.str += '\n'; break;
            case 't': j.str += '\t'; break;
            case 'r': j.str += '\r'; break;
            case 'b': j.str += '\b'; break;
            case 'f': j.str += '\f'; break;
            case 'u': {
              if (end - p < 4) { ok = false; break; }
              unsigned v = 0;
              for (int i = 0; i < 4; i++) {
                char h = p[i];
                v = v * 16 + (h >= '0' && h <= '9' ? h - '0' : h >= 'a' && h <= 'f' ? h - 'a' + 10 : h >= 'A' && h <= 'F' ? h - 'A' + 10 : 0);
              }
              p += 4;
              if (v < 0x80) j.str += (char)v;
              else if (v < 0x800) { j.str += (char)(0xC0 | (v >> 6)); j.str += (char)(0x80 | (v & 0x3F)); }
              else { j.str += (char)(0xE0 | (v >> 12)); j.str += (char)(0x80 | ((v >> 6) & 0x3F)); j.str += (char)(0x80 | (v & 0x3F)); }
              break;
            }
            default: j.str += c;  // \" \\ \/
          }
        } else {
          j.str += *p++;
        }
      }
      if (p >= end) ok = false; else p++;
    } else if (lit("true")) { j.kind = Json::Bool; j.b = true; }
    else if (lit("false")) { j.kind = Json::Bool; }
    else if (lit("null")) { }
    else {
      j.kind = Json::Num;
      const char* s = p;
      while (p < end && (*p == '-' || *p == '+' || *p == '.' || *p == 'e' || *p == 'E' || (*p >= '0' && *p <= '9'))) p++;
      if (p == s || p - s > 40) { ok = false; }
      else j.num = strtod(std::string(s, p).c_str(), nullptr);
    }
    depth--;
    return j;
  }
};

static bool parse_json(const std::string& s, Json* out) {
  JsonParser jp{s.data(), s.data() + s.size()};
  *out = jp.value();
  jp.ws();
  return jp.ok && jp.p == jp.end;
}

// ---- the Lance-v2 container ----
struct V2File {
  const uint8_t* blob = nullptr;
  size_t size = 0;
  const char* what = "";
  std::vector<std::pair<uint64_t, uint64_t>> cols, gbufs;
  struct Field { std::string name, ltype; int64_t parent = 0; };
  std::vector<Field> fields;
  std::map<std::string, std::string> meta;

  bool slice(uint64_t pos, uint64_t len, const uint8_t** p) const {
    if (pos > size || len > size - pos) return false;
    *p = blob + pos;
    return true;
  }
  int open(const uint8_t* b, size_t n, const char* w) {
    blob = b; size = n; what = w;
    if (n < 40) return set_err(LB200_EINVAL, "%s: does not have sufficient data, len: %zu", w, n);
    if (memcmp(b + n - 4, "LANC", 4) != 0) return set_err(LB200_EINVAL, "%s: not a Lance file: magic number mismatch", w);
    uint64_t cm, cmo, gbo;
    uint32_t ngb, ncol;
    uint16_t major, minor;
    memcpy(&cm, b + n - 40, 8); memcpy(&cmo, b + n - 32, 8); memcpy(&gbo, b + n - 24, 8);
    memcpy(&ngb, b + n - 16, 4); memcpy(&ncol, b + n - 12, 4); memcpy(&major, b + n - 8, 2); memcpy(&minor, b + n - 6, 2);
    if (!((major == 0 && minor == 3) || (major == 2 && minor == 0)))
      return set_err(LB200_EUNSUPPORTED, "%s: file version %u.%u is not the Lance 2.0 container of a v3 vector index", w, major, minor);
    const uint64_t end = n - 40;
    // every offset comes from the file: ordered, inside the file, and the table sizes must match the counts exactly
    if (!(cm <= cmo && cmo <= gbo && gbo <= end) || (gbo - cmo) != 16ull * ncol || (end - gbo) != 16ull * ngb)
      return set_err(LB200_EINVAL, "%s: inconsistent footer offsets", w);
    if (ngb < 1) return set_err(LB200_EINVAL, "%s: no global buffers (the schema lives in buffer 0)", w);
    cols.resize(ncol);
    gbufs.resize(ngb);
    for (uint32_t i = 0; i < ncol; i++) { memcpy(&cols[i].first, b + cmo + 16ull * i, 8); memcpy(&cols[i].second, b + cmo + 16ull * i + 8, 8); }
    for (uint32_t i = 0; i < ngb; i++) { memcpy(&gbufs[i].first, b + gbo + 16ull * i, 8); memcpy(&gbufs[i].second, b + gbo + 16ull * i + 8, 8); }
    const uint8_t* d;
    if (!slice(gbufs[0].first, gbufs[0].second, &d)) return set_err(LB200_EINVAL, "%s: global buffer runs past the end of the file", w);
    // pb::FileDescriptor { Schema schema = 1 { repeated Field fields = 1; map<string,bytes> metadata = 5 }; length = 2 }
    PbReader r(d, gbufs[0].second);
    uint32_t fno, wt;
    uint64_t val;
    const uint8_t* data;
    size_t len;
    while (r.next(fno, wt, val, data, len)) {
      if (fno != 1 || wt != 2) continue;
      PbReader rs(data, len);
      uint32_t f2, w2;
      uint64_t v2;
      const uint8_t* d2;
      size_t l2;
      while (rs.next(f2, w2, v2, d2, l2)) {
        if (f2 == 1 && w2 == 2) {
          Field f;
          PbReader rf(d2, l2);
          uint32_t f3, w3;
          uint64_t v3;
          const uint8_t* d3;
          size_t l3;
          while (rf.next(f3, w3, v3, d3, l3)) {
            if (f3 == 2 && w3 == 2) f.name.assign((const char*)d3, l3);
            else if (f3 == 4) f.parent = (int64_t)v3;
            else if (f3 == 5 && w3 == 2) f.ltype.assign((const char*)d3, l3);
          }
          if (!rf.ok) return set_err(LB200_EINVAL, "%s: malformed schema field", w);
          fields.push_back(std::move(f));
        } else if (f2 == 5 && w2 == 2) {
          std::string k, v;
          PbReader rm(d2, l2);
          uint32_t f3, w3;
          uint64_t v3;
          const uint8_t* d3;
          size_t l3;
          while (rm.next(f3, w3, v3, d3, l3)) {
            if (f3 == 1 && w3 == 2) k.assign((const char*)d3, l3);
            else if (f3 == 2 && w3 == 2) v.assign((const char*)d3, l3);
          }
          if (!rm.ok) return set_err(LB200_EINVAL, "%s: malformed schema metadata", w);
          meta[k] = v;
        }
      }
      if (!rs.ok) return set_err(LB200_EINVAL, "%s: malformed schema", w);
    }
    if (!r.ok) return set_err(LB200_EINVAL, "%s: malformed file descriptor", w);
    return 0;
  }
  int global_buffer(uint64_t idx, const uint8_t** p, uint64_t* len) const {
    if (idx >= gbufs.size()) return set_err(LB200_EINVAL, "%s: global buffer %llu does not exist", what, (unsigned long long)idx);
    if (!slice(gbufs[idx].first, gbufs[idx].second, p)) return set_err(LB200_EINVAL, "%s: global buffer runs past the end of the file", what);
    *len = gbufs[idx].second;
    return 0;
  }
  int column_index(const char* name, uint32_t* out) const {
    uint32_t i = 0;
    for (auto& f : fields) {
      if (f.parent >= 0) continue;  // parent_id -1 = top level; primitive / fixed-size-list fields take one column each
      if (f.name == name) { *out = i; return 0; }
      i++;
    }
    return set_err(LB200_EINVAL, "%s: column %s not found", what, name);
  }
};

static bool pb_sub(const uint8_t* b, size_t n, uint32_t want, const uint8_t** out, size_t* outlen) {
  PbReader r(b, n);
  uint32_t fno, wt;
  uint64_t val;
  const uint8_t* data;
  size_t len;
  while (r.next(fno, wt, val, data, len))
    if (fno == want && wt == 2) { *out = data; *outlen = len; return true; }
  return false;
}

// ArrayEncoding -> buffer index of its Flat leaf through Nullable{no_nulls}; bits per value checked, no compression
static int v3_flat(const uint8_t* enc, size_t n, uint64_t bits, uint64_t* bufidx) {
  const uint8_t* s;
  size_t l;
  if (pb_sub(enc, n, 2, &s, &l)) {  // nullable
    const uint8_t *a, *b2;
    size_t la, lb;
    if (pb_sub(s, l, 2, &a, &la) || pb_sub(s, l, 3, &a, &la)) return set_err(LB200_EUNSUPPORTED, "null values are not supported in index columns");
    if (!pb_sub(s, l, 1, &a, &la) || !pb_sub(a, la, 1, &b2, &lb)) return set_err(LB200_EINVAL, "malformed Nullable page encoding");
    enc = b2;
    n = lb;
  }
  if (!pb_sub(enc, n, 1, &s, &l)) return set_err(LB200_EUNSUPPORTED, "unsupported page encoding (expected Flat)");
  PbReader r(s, l);
  uint32_t fno, wt;
  uint64_t val, got = 0;
  const uint8_t* data;
  size_t len;
  bool compressed = false;
  *bufidx = 0;
  while (r.next(fno, wt, val, data, len)) {
    if (fno == 1) got = val;
    else if (fno == 2 && wt == 2) {
      PbReader rb(data, len);
      uint32_t f2, w2;
      uint64_t v2;
      const uint8_t* d2;
      size_t l2;
      while (rb.next(f2, w2, v2, d2, l2))
        if (f2 == 1) *bufidx = v2;
    } else if (fno == 3) compressed = true;
  }
  if (!r.ok || got != bits || compressed)
    return set_err(LB200_EUNSUPPORTED, "unsupported Flat page encoding (bits %llu, compressed %d)", (unsigned long long)got, (int)compressed);
  return 0;
}

// all pages of one column, concatenated: rows * bytes_per_row bytes appended to out
static int v3_column(const V2File& f, uint32_t col, uint64_t bytes_per_row, uint64_t fsl_dim, uint64_t max_rows,
                     std::vector<uint8_t>& out, uint64_t* rows_out) {
  if (col >= f.cols.size()) return set_err(LB200_EINVAL, "%s: column %u has no metadata", f.what, col);
  const uint8_t* cm;
  if (!f.slice(f.cols[col].first, f.cols[col].second, &cm)) return set_err(LB200_EINVAL, "%s: column metadata runs past the end of the file", f.what);
  PbReader r(cm, f.cols[col].second);
  uint32_t fno, wt;
  uint64_t val, rows = 0;
  const uint8_t* page;
  size_t plen;
  while (r.next(fno, wt, val, page, plen)) {
    if (fno != 2 || wt != 2) continue;
    std::vector<uint64_t> offs, sizes;
    uint64_t length = 0;
    const uint8_t* enc = nullptr;
    size_t elen = 0;
    bool ok = true;
    PbReader rp(page, plen);
    uint32_t f2, w2;
    uint64_t v2;
    const uint8_t* d2;
    size_t l2;
    while (rp.next(f2, w2, v2, d2, l2)) {
      if (f2 == 1) pb_packed_u64(w2, v2, d2, l2, offs, ok);
      else if (f2 == 2) pb_packed_u64(w2, v2, d2, l2, sizes, ok);
      else if (f2 == 3) length = v2;
      else if (f2 == 4 && w2 == 2) { enc = d2; elen = l2; }
    }
    if (!rp.ok || !ok || !enc) return set_err(LB200_EINVAL, "%s: malformed page", f.what);
    const uint8_t *direct, *any, *payload = nullptr, *url = nullptr;
    size_t ldirect, lany, lpayload = 0, lurl = 0;
    if (!pb_sub(enc, elen, 2, &direct, &ldirect) || !pb_sub(direct, ldirect, 1, &any, &lany))
      return set_err(LB200_EUNSUPPORTED, "%s: page without a direct encoding", f.what);
    pb_sub(any, lany, 1, &url, &lurl);
    if (!pb_sub(any, lany, 2, &payload, &lpayload)) { payload = any; lpayload = 0; }
    static const char kUrl[] = "lance.encodings.ArrayEncoding";
    if (lurl < sizeof(kUrl) - 1 || memcmp(url + lurl - (sizeof(kUrl) - 1), kUrl, sizeof(kUrl) - 1) != 0)
      return set_err(LB200_EUNSUPPORTED, "%s: page encoding is not a format-2.0 ArrayEncoding", f.what);
    uint64_t bufidx = 0;
    int rc;
    if (fsl_dim) {
      const uint8_t *inner = payload, *a, *b2, *fsl;
      size_t linner = lpayload, la, lb, lfsl;
      if (pb_sub(payload, lpayload, 2, &a, &la)) {
        if (!pb_sub(a, la, 1, &b2, &lb) || !pb_sub(b2, lb, 1, &inner, &linner)) return set_err(LB200_EUNSUPPORTED, "%s: nullable FixedSizeList pages are not supported", f.what);
      }
      if (!pb_sub(inner, linner, 3, &fsl, &lfsl)) return set_err(LB200_EUNSUPPORTED, "%s: expected a FixedSizeList page encoding", f.what);
      PbReader rf(fsl, lfsl);
      uint64_t dim = 0;
      const uint8_t* items = nullptr;
      size_t litems = 0;
      while (rf.next(f2, w2, v2, d2, l2)) {
        if (f2 == 1) dim = v2;
        else if (f2 == 2 && w2 == 2) { items = d2; litems = l2; }
      }
      if (!rf.ok || dim != fsl_dim || !items) return set_err(LB200_EINVAL, "%s: FixedSizeList dimension %llu != %llu", f.what, (unsigned long long)dim, (unsigned long long)fsl_dim);
      rc = v3_flat(items, litems, 8, &bufidx);
    } else {
      rc = v3_flat(payload, lpayload, 8 * bytes_per_row, &bufidx);
    }
    if (rc) return rc;
    if (bufidx >= offs.size() || offs.size() != sizes.size()) return set_err(LB200_EINVAL, "%s: page buffer %llu missing", f.what, (unsigned long long)bufidx);
    // sizes come from the file: rows bounded by the caller's total BEFORE anything is appended
    if (length > max_rows - rows) return set_err(LB200_EINVAL, "%s: column holds more rows than the IVF model", f.what);
    const uint64_t need = length * bytes_per_row;
    const uint8_t* src;
    if (sizes[bufidx] < need || !f.slice(offs[bufidx], need, &src)) return set_err(LB200_EINVAL, "%s: page buffer too small or past the end of the file", f.what);
    out.insert(out.end(), src, src + need);
    rows += length;
  }
  if (!r.ok) return set_err(LB200_EINVAL, "%s: malformed column metadata", f.what);
  *rows_out = rows;
  return 0;
}

static int parse_ivf_pb(const uint8_t* b, size_t n, std::vector<float>* cents, uint64_t cshape[2], std::vector<uint64_t>* lengths) {
  PbReader ri(b, n);
  uint32_t a, w;
  uint64_t c;
  const uint8_t* d;
  size_t l;
  bool ok = true;
  std::vector<uint64_t> offsets;
  while (ri.next(a, w, c, d, l)) {
    if (a == 2) pb_packed_u64(w, c, d, l, offsets, ok);
    else if (a == 3) pb_packed_u64(w, c, d, l, *lengths, ok);
    else if (a == 4 && w == 2 && cents) { int rc = pb_tensor(d, l, *cents, cshape); if (rc) return rc; }
  }
  if (!ri.ok || !ok) return set_err(LB200_EINVAL, "malformed IVF model");
  return 0;
}

static int parse_v3_index(const uint8_t* ib, size_t in, const uint8_t* ab, size_t an, lb200_index_file* f) {
  V2File idx, aux;
  int rc = idx.open(ib, in, "index.idx");
  if (rc) return rc;
  if ((rc = aux.open(ab, an, "auxiliary.idx"))) return rc;
  auto need = [&](const V2File& v, const char* k, std::string* out) {
    auto it = v.meta.find(k);
    if (it == v.meta.end()) return set_err(LB200_EINVAL, "%s: schema metadata %s not found", v.what, k);
    *out = it->second;
    return 0;
  };
  std::string s;
  Json j;
  if ((rc = need(idx, "lance:index", &s))) return rc;
  if (!parse_json(s, &j) || j.kind != Json::Obj || !j.get("type") || !j.get("distance_type"))
    return set_err(LB200_EINVAL, "index.idx: malformed lance:index metadata");
  if (j.get("type")->str != "IVF_PQ") return set_err(LB200_EUNSUPPORTED, "index type %s is not IVF_PQ", j.get("type")->str.c_str());
  std::string metric = j.get("distance_type")->str;
  for (auto& ch : metric) ch = (char)tolower(ch);
  if (metric == "l2" || metric == "euclidean") f->metric = LB200_L2;
  else if (metric == "cosine") f->metric = LB200_COSINE;
  else if (metric == "dot") f->metric = LB200_DOT;
  else return set_err(LB200_EINVAL, "Metric type '%s' is not supported", metric.c_str());
  auto mit = aux.meta.find("distance_type");
  if (mit != aux.meta.end()) {
    std::string m2 = mit->second;
    for (auto& ch : m2) ch = (char)tolower(ch);
    if (m2 != metric) return set_err(LB200_EINVAL, "index.idx and auxiliary.idx disagree on the distance type");
  }
  // IVF models: centroids from index.idx, partition lengths from auxiliary.idx
  const uint8_t* gb;
  uint64_t gl;
  uint64_t cshape[2] = {0, 0};
  std::vector<uint64_t> ilen, lengths;
  if ((rc = need(idx, "lance:ivf", &s))) return rc;
  if ((rc = idx.global_buffer(strtoull(s.c_str(), nullptr, 10), &gb, &gl))) return rc;
  if ((rc = parse_ivf_pb(gb, gl, &f->centroids, cshape, &ilen))) return rc;
  if ((rc = need(aux, "lance:ivf", &s))) return rc;
  if ((rc = aux.global_buffer(strtoull(s.c_str(), nullptr, 10), &gb, &gl))) return rc;
  if ((rc = parse_ivf_pb(gb, gl, nullptr, nullptr, &lengths))) return rc;
  if (f->centroids.empty() || !cshape[0] || !cshape[1]) return set_err(LB200_EINVAL, "index.idx: IVF model has no centroids");
  if (cshape[0] > 0xffffffffull || cshape[1] > 0xffffffffull) return set_err(LB200_EINVAL, "index.idx: centroid tensor too large");
  f->nlist = (uint32_t)cshape[0];
  f->dim = (uint32_t)cshape[1];
  if (lengths.size() != f->nlist) return set_err(LB200_EINVAL, "auxiliary.idx: %zu partitions, index.idx has %u", lengths.size(), f->nlist);
  // PQ metadata
  if ((rc = need(aux, "storage_metadata", &s))) return rc;
  Json sm, pq;
  if (!parse_json(s, &sm) || sm.kind != Json::Arr || sm.arr.empty() || sm.arr[0].kind != Json::Str || !parse_json(sm.arr[0].str, &pq) ||
      pq.kind != Json::Obj)
    return set_err(LB200_EINVAL, "auxiliary.idx: malformed storage_metadata");
  const Json *jm = pq.get("num_sub_vectors"), *jb = pq.get("num_bits"), *jd = pq.get("dimension"), *jt = pq.get("codebook_tensor"),
             *jtr = pq.get("transposed");
  if (!jm || !jb || !jd || !jt || jt->kind != Json::Arr) return set_err(LB200_EINVAL, "auxiliary.idx: incomplete PQ metadata");
  if (jm->num < 1 || jm->num > 65536 || jd->num != (double)f->dim) return set_err(LB200_EINVAL, "PQ metadata does not match the IVF dimension");
  f->M = (uint32_t)jm->num;
  f->num_bits = (uint32_t)jb->num;
  if (f->num_bits != 8) return set_err(LB200_EUNSUPPORTED, "num_bits %u: only 8-bit PQ is on the B200 path", f->num_bits);
  if (f->dim % f->M) return set_err(LB200_EINVAL, "PQ stage: dimension %u is not divisible by num_sub_vectors %u", f->dim, f->M);
  std::vector<uint8_t> tb(jt->arr.size());
  for (size_t i = 0; i < tb.size(); i++) {
    const Json& e = jt->arr[i];
    if (e.kind != Json::Num || e.num < 0 || e.num > 255) return set_err(LB200_EINVAL, "auxiliary.idx: codebook tensor is not a byte array");
    tb[i] = (uint8_t)e.num;
  }
  uint64_t bshape[2];
  if ((rc = pb_tensor(tb.data(), tb.size(), f->codebook, bshape))) return rc;
  if (f->codebook.size() != (size_t)256 * f->dim) return set_err(LB200_EINVAL, "PQ codebook has %zu values, expected 256x%u", f->codebook.size(), f->dim);
  const bool transposed = jtr && jtr->kind == Json::Bool && jtr->b;
  // rows: validated against the file size before anything is allocated from them
  const uint32_t M = f->M;
  const uint64_t max_rows = an / (M + 8ull);
  f->part_offsets.assign(f->nlist + 1, 0);
  for (uint32_t p = 0; p < f->nlist; p++) {
    if (lengths[p] > max_rows - f->part_offsets[p]) return set_err(LB200_EINVAL, "IVF model: partition lengths add up to more rows than the file can hold");
    f->part_offsets[p + 1] = f->part_offsets[p] + lengths[p];
  }
  const uint64_t N = f->part_offsets[f->nlist];
  uint32_t rid_col, pq_col;
  if ((rc = aux.column_index("_rowid", &rid_col)) || (rc = aux.column_index("__pq_code", &pq_col))) return rc;
  std::vector<uint8_t> rid_bytes;
  uint64_t n1 = 0, n2 = 0;
  if ((rc = v3_column(aux, rid_col, 8, 0, N, rid_bytes, &n1))) return rc;
  if ((rc = v3_column(aux, pq_col, M, M, N, f->codes, &n2))) return rc;
  if (n1 != N || n2 != N)
    return set_err(LB200_EINVAL, "auxiliary.idx holds %llu/%llu rows, the IVF model %llu", (unsigned long long)n1, (unsigned long long)n2, (unsigned long long)N);
  f->row_ids.resize(N);
  if (N) memcpy(f->row_ids.data(), rid_bytes.data(), N * 8);
  if (!transposed) {  // row-major per partition -> [M][n_p] (transpose(), pq/storage.rs:357-377)
    std::vector<uint8_t> t;
    for (uint32_t p = 0; p < f->nlist; p++) {
      const uint64_t o = f->part_offsets[p], n = lengths[p];
      t.assign(f->codes.begin() + o * M, f->codes.begin() + (o + n) * M);
      uint8_t* dst = f->codes.data() + o * M;
      for (uint64_t i = 0; i < n; i++)
        for (uint32_t m = 0; m < M; m++) dst[(size_t)m * n + i] = t[i * M + m];
    }
  }
  f->lengths.resize(f->nlist);
  for (uint32_t p = 0; p < f->nlist; p++) f->lengths[p] = (uint32_t)lengths[p];
  f->spec_version = 3;
  f->column = "";
  f->name = "";
  return 0;
}

static int read_whole_file(const char* path, std::vector<uint8_t>* out) {
  std::ifstream in(path, std::ios::binary | std::ios::ate);
  if (!in) return set_err(LB200_EINVAL, "cannot open index file %s", path);
  const std::streamsize n = in.tellg();
  out->resize((size_t)std::max<std::streamsize>(n, 0));
  in.seekg(0);
  if (n > 0 && !in.read((char*)out->data(), n)) return set_err(LB200_EINVAL, "cannot read index file %s", path);
  return 0;
}

}  // namespace

extern "C" int32_t lb200_index_file_open_v3_bytes(const void* index_idx, uint64_t index_size, const void* auxiliary_idx,
                                                  uint64_t auxiliary_size, lb200_index_file** out) {
  if (!out) return set_err(LB200_EINVAL, "null output");
  *out = nullptr;
  if ((!index_idx && index_size) || (!auxiliary_idx && auxiliary_size)) return set_err(LB200_EINVAL, "null argument");
  lb200_index_file* f = nullptr;
  try {
    f = new lb200_index_file();
    int rc = parse_v3_index((const uint8_t*)index_idx, (size_t)index_size, (const uint8_t*)auxiliary_idx, (size_t)auxiliary_size, f);
    if (rc) { delete f; return rc; }
  } catch (const std::bad_alloc&) {
    delete f;
    return set_err(LB200_ENOMEM, "out of host memory while parsing the index files");
  } catch (...) {
    delete f;
    return set_err(LB200_EINVAL, "malformed index files");
  }
  *out = f;
  return 0;
}

extern "C" int32_t lb200_index_file_open_v3(const char* index_idx_path, const char* auxiliary_idx_path, lb200_index_file** out) {
  if (!out) return set_err(LB200_EINVAL, "null output");
  *out = nullptr;
  if (!index_idx_path || !auxiliary_idx_path) return set_err(LB200_EINVAL, "null path");
  try {
    std::vector<uint8_t> a, b;
    int rc = read_whole_file(index_idx_path, &a);
    if (rc) return rc;
    if ((rc = read_whole_file(auxiliary_idx_path, &b))) return rc;
    return lb200_index_file_open_v3_bytes(a.data(), a.size(), b.data(), b.size(), out);
  } catch (const std::bad_alloc&) {
    return set_err(LB200_ENOMEM, "out of host memory while reading the index files");
  } catch (...) {
    return set_err(LB200_EINVAL, "cannot read the index files");
  }
}

extern "C" int32_t lb200_index_open_v3(int32_t device_id, const char* index_idx_path, const char* auxiliary_idx_path,
                                       const float* raw_vectors_or_null, lb200_index** out) {
  if (!out) return set_err(LB200_EINVAL, "null output");
  *out = nullptr;
  lb200_index_file* f = nullptr;
  int rc = lb200_index_file_open_v3(index_idx_path, auxiliary_idx_path, &f);
  if (rc) return rc;
  rc = lb200_index_create(device_id, f->metric, f->dim, f->nlist, f->centroids.data(), f->M, f->num_bits, f->codebook.data(),
                          f->part_offsets.data(), f->codes.data(), f->row_ids.data(), raw_vectors_or_null, out);
  delete f;
  return rc;
}
