// Host-side readers of the two on-disk formats that feed the stepping path:
//
//   * PNG heightmaps -- ode::HeightField's image constructor (ode/heightfield.hh:42-59) goes through
//     osgDB::readImageFile and keeps the FIRST BYTE of every pixel (`*heightimage->data(c, r)`), scaled by 0.3/255,
//     with OSG's bottom-up rows flipped back, so row 0 of the height array is the top row of the picture.  OSG and
//     libpng are not part of this engine: decode_png() below is a self-contained decoder (zlib for the inflate step)
//     for what the reference's maps are -- non-interlaced PNG, bit depth 1/2/4/8, grey / grey+alpha / RGB / RGBA /
//     palette -- and fails loudly on everything else.
//   * model-tree YAML -- Linear_model_tree's constructor (scripts/ModelTree/cpp/model_tree.cc:11-45) reads, through
//     yaml-cpp, a map  node id -> { info, terminal, "model params" | "split params" }  written by PyYAML
//     (scripts/ModelTree/model_tree.py:514-520).  parse_model_tree() reads that subset of YAML (block or flow
//     sequences, comments, quoted or plain scalars) and flattens the tree in the constructor's own order: node ids are
//     assigned depth-first, child 0 before child 1, exactly as _build_tree_recursively/_new_node do.
//
//   * TensorFlow SavedModel variables -- TF_model's constructor (scripts/MachineLearning/tf_cpp_binding/
//     tf_cpp_binding.cc:14-66) hands the directory written by `td3.actor.save( session_dir + '/actor' )`
//     (scripts/rover_training_1.py:90,121) to TF_LoadSessionFromSavedModel.  TensorFlow is not part of this engine:
//     read_saved_model_actor() reads the checkpoint inside that directory directly -- `variables/variables.index`
//     (a LevelDB-format sorted table, uncompressed blocks, key -> BundleEntryProto) and the
//     `variables/variables.data-XXXXX-of-YYYYY` shards -- and returns the three Dense layers of
//     scripts/rover_training_1.py:26-35 as the blob tfode_batch_load_policy takes.  FORMAT UNPINNED: no SavedModel
//     ships with the reference and TensorFlow is absent on this box, so the layout is restated from TensorFlow's
//     published tensor_bundle / table format and tested only against a writer of the same restatement
//     (tests/test_formats.py); every deviation fails loudly (magic, compression, dtype, shape, CRC32C).
//
// Nothing here steps an environment; this is format plumbing for tfode_capi.cu.
#pragma once
#include <zlib.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace tfode_fmt {

// ------------------------------------------------------------------------------------------------ PNG
inline uint32_t be32(const unsigned char* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3]; }

// first_byte[row*ncol + col] = first byte of pixel (col, row) after libpng-style expansion to 8 bits (what OSG's reader hands
// to the reference); row 0 = top row of the picture.
inline void decode_png(const unsigned char* d, size_t n, std::vector<unsigned char>& first_byte, int& nrow, int& ncol) {
  static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  if (n < 8 || std::memcmp(d, sig, 8) != 0) throw std::runtime_error("not a PNG file (bad signature)");
  size_t p = 8;
  uint32_t W = 0, H = 0; int depth = 0, ctype = -1, interlace = 0;
  std::vector<unsigned char> idat, plte;
  bool seen_ihdr = false, seen_iend = false;
  while (p + 12 <= n && !seen_iend) {
    const uint32_t len = be32(d + p);
    if (len > n || p + 12 + (size_t)len > n) throw std::runtime_error("PNG chunk runs past the end of the file");
    const unsigned char* type = d + p + 4; const unsigned char* body = d + p + 8;
    const uint32_t crc = be32(body + len);
    if ((uint32_t)crc32(crc32(0L, Z_NULL, 0), type, len + 4) != crc) throw std::runtime_error("PNG chunk CRC mismatch");
    if (!std::memcmp(type, "IHDR", 4)) {
      if (len != 13) throw std::runtime_error("bad IHDR length");
      W = be32(body); H = be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12];
      if (body[10] != 0 || body[11] != 0) throw std::runtime_error("unknown PNG compression/filter method");
      seen_ihdr = true;
    } else if (!seen_ihdr) throw std::runtime_error("PNG does not start with IHDR");
    else if (!std::memcmp(type, "PLTE", 4)) plte.assign(body, body + len);
    else if (!std::memcmp(type, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
    else if (!std::memcmp(type, "IEND", 4)) seen_iend = true;
    else if (!(type[0] & 0x20)) throw std::runtime_error("unknown critical PNG chunk");
    p += 12 + (size_t)len;
  }
  if (!seen_ihdr || !seen_iend) throw std::runtime_error("truncated PNG (IHDR/IEND missing)");
  if (W == 0 || H == 0 || W > 16384 || H > 16384) throw std::runtime_error("PNG dimensions out of range");
  if (interlace != 0) throw std::runtime_error("interlaced PNG heightmaps are not supported");
  int channels;
  switch (ctype) { case 0: channels = 1; break; case 2: channels = 3; break; case 3: channels = 1; break; case 4: channels = 2; break; case 6: channels = 4; break;
    default: throw std::runtime_error("bad PNG colour type"); }
  const bool depth_ok = (ctype == 0 && (depth == 1 || depth == 2 || depth == 4 || depth == 8)) || (ctype == 3 && (depth == 1 || depth == 2 || depth == 4 || depth == 8)) ||
                        ((ctype == 2 || ctype == 4 || ctype == 6) && depth == 8);
  if (!depth_ok) throw std::runtime_error(depth == 16 ? "16-bit PNG heightmaps are not supported (the reference keeps one byte per pixel)" : "bad PNG bit depth");
  if (ctype == 3 && plte.empty()) throw std::runtime_error("palette PNG without PLTE");
  const size_t stride = ((size_t)W * channels * depth + 7) / 8, bpp = (size_t)std::max(1, channels * depth / 8);
  std::vector<unsigned char> raw((stride + 1) * H);
  uLongf out_len = (uLongf)raw.size();
  const int zr = uncompress(raw.data(), &out_len, idat.data(), (uLong)idat.size());
  if (zr != Z_OK || out_len != raw.size()) throw std::runtime_error("PNG image data does not inflate to width x height");
  // undo the per-scanline filters in place (PNG spec 9.2)
  std::vector<unsigned char> zero(stride, 0);
  for (uint32_t y = 0; y < H; y++) {
    unsigned char* line = raw.data() + (stride + 1) * y;
    const int f = line[0];
    unsigned char* cur = line + 1;
    const unsigned char* up = y ? line - stride : zero.data();
    if (f < 0 || f > 4) throw std::runtime_error("bad PNG filter type");
    for (size_t i = 0; i < stride; i++) {
      const int a = i >= bpp ? cur[i - bpp] : 0, b = up[i], c = i >= bpp ? up[i - bpp] : 0;
      int pred = 0;
      if (f == 1) pred = a; else if (f == 2) pred = b; else if (f == 3) pred = (a + b) >> 1;
      else if (f == 4) { const int q = a + b - c, pa = std::abs(q - a), pb = std::abs(q - b), pc = std::abs(q - c); pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); }
      cur[i] = (unsigned char)(cur[i] + pred);
    }
  }
  nrow = (int)H; ncol = (int)W;
  first_byte.resize((size_t)W * H);
  for (uint32_t y = 0; y < H; y++) {
    const unsigned char* cur = raw.data() + (stride + 1) * y + 1;
    for (uint32_t x = 0; x < W; x++) {
      unsigned v;
      if (depth == 8) v = cur[(size_t)x * channels];
      else { const unsigned bit = x * depth; v = (cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1u << depth) - 1); }
      if (ctype == 3) { if ((size_t)v * 3 >= plte.size()) throw std::runtime_error("palette index out of range"); v = plte[(size_t)v * 3]; }
      else if (depth < 8) v = v * 255u / ((1u << depth) - 1);   // png_set_expand_gray_1_2_4_to_8
      first_byte[(size_t)y * W + x] = (unsigned char)v;
    }
  }
}

inline std::vector<unsigned char> read_file(const char* path) {
  FILE* f = std::fopen(path, "rb");
  if (!f) throw std::runtime_error(std::string("Can't open ") + path);   // message of ode/heightfield.hh:48
  std::vector<unsigned char> buf;
  unsigned char tmp[65536]; size_t k;
  while ((k = std::fread(tmp, 1, sizeof(tmp), f)) > 0) buf.insert(buf.end(), tmp, tmp + k);
  std::fclose(f);
  return buf;
}

// ------------------------------------------------------------------------------------------------ model-tree YAML
struct TreeNode { int32_t id, terminal, child[2], param_off, n_params; };
struct FlatTree { std::vector<TreeNode> nodes; std::vector<double> params; };

struct YamlNodeEntry { bool has_terminal = false, terminal = false, has_model = false, has_split = false; std::vector<double> model, split; };

inline std::string trim(const std::string& s) {
  size_t a = 0, b = s.size();
  while (a < b && (s[a] == ' ' || s[a] == '\t' || s[a] == '\r')) a++;
  while (b > a && (s[b - 1] == ' ' || s[b - 1] == '\t' || s[b - 1] == '\r')) b--;
  return s.substr(a, b - a);
}
inline std::string unquote(const std::string& s) {
  if (s.size() >= 2 && ((s.front() == '"' && s.back() == '"') || (s.front() == '\'' && s.back() == '\''))) return s.substr(1, s.size() - 2);
  return s;
}
inline std::string strip_comment(const std::string& s) {   // '#' starts a comment at the line start or after a blank, outside quotes
  char q = 0;
  for (size_t i = 0; i < s.size(); i++) {
    if (q) { if (s[i] == q) q = 0; }
    else if (s[i] == '"' || s[i] == '\'') q = s[i];
    else if (s[i] == '#' && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) return s.substr(0, i);
  }
  return s;
}
inline double parse_number(const std::string& raw, int line_no) {
  std::string s = unquote(trim(raw));
  if (s == ".inf" || s == "+.inf" || s == ".Inf") return INFINITY;
  if (s == "-.inf" || s == "-.Inf") return -INFINITY;
  if (s == ".nan" || s == ".NaN") return NAN;
  char* end = nullptr;
  const double v = std::strtod(s.c_str(), &end);
  if (s.empty() || end != s.c_str() + s.size()) throw std::runtime_error("model tree YAML line " + std::to_string(line_no) + ": '" + s + "' is not a number");
  return v;
}
inline bool parse_bool(const std::string& raw, int line_no) {
  std::string s = unquote(trim(raw));
  for (auto& ch : s) ch = (char)std::tolower((unsigned char)ch);
  if (s == "true" || s == "yes" || s == "on") return true;
  if (s == "false" || s == "no" || s == "off") return false;
  throw std::runtime_error("model tree YAML line " + std::to_string(line_no) + ": '" + s + "' is not a boolean");
}
inline void parse_flow_seq(const std::string& raw, std::vector<double>& out, int line_no) {
  const std::string s = trim(raw);
  if (s.size() < 2 || s.front() != '[' || s.back() != ']') throw std::runtime_error("model tree YAML line " + std::to_string(line_no) + ": expected a [ ... ] sequence");
  const std::string body = s.substr(1, s.size() - 2);
  if (trim(body).empty()) return;
  size_t a = 0;
  while (a <= body.size()) {
    size_t c = body.find(',', a);
    if (c == std::string::npos) c = body.size();
    out.push_back(parse_number(body.substr(a, c - a), line_no));
    a = c + 1;
  }
}

// text -> { node id -> entry }
inline std::map<int, YamlNodeEntry> parse_tree_yaml(const char* text, size_t len) {
  std::map<int, YamlNodeEntry> nodes;
  YamlNodeEntry* cur = nullptr;
  std::vector<double>* seq = nullptr;       // open block sequence
  std::string pending_flow; std::vector<double>* flow_dst = nullptr; int flow_line = 0;   // flow sequence spanning lines
  size_t pos = 0; int line_no = 0;
  while (pos <= len) {
    size_t e = pos;
    while (e < len && text[e] != '\n') e++;
    std::string line = strip_comment(std::string(text + pos, e - pos));
    pos = e + 1; line_no++;
    if (flow_dst) {
      pending_flow += " " + trim(line);
      if (pending_flow.find(']') != std::string::npos) { parse_flow_seq(pending_flow, *flow_dst, flow_line); flow_dst = nullptr; pending_flow.clear(); }
      continue;
    }
    const std::string t = trim(line);
    if (t.empty() || t == "---" || t == "...") continue;
    size_t indent = 0;
    while (indent < line.size() && line[indent] == ' ') indent++;
    if (indent < line.size() && line[indent] == '\t') throw std::runtime_error("model tree YAML line " + std::to_string(line_no) + ": tab indentation");
    if (t[0] == '-' && (t.size() == 1 || t[1] == ' ')) {   // block sequence item
      if (!seq) throw std::runtime_error("model tree YAML line " + std::to_string(line_no) + ": sequence item outside 'model params' / 'split params'");
      seq->push_back(parse_number(t.substr(1), line_no));
      continue;
    }
    // key: value (the key may be quoted and may contain blanks: "model params")
    size_t colon = std::string::npos; char q = 0;
    for (size_t i = 0; i < t.size(); i++) {
      if (q) { if (t[i] == q) q = 0; }
      else if (t[i] == '"' || t[i] == '\'') q = t[i];
      else if (t[i] == ':' && (i + 1 == t.size() || t[i + 1] == ' ')) { colon = i; break; }
    }
    if (colon == std::string::npos) throw std::runtime_error("model tree YAML line " + std::to_string(line_no) + ": expected 'key: value'");
    const std::string key = unquote(trim(t.substr(0, colon))), val = trim(t.substr(colon + 1));
    seq = nullptr;
    if (indent == 0) {   // node id
      char* end = nullptr;
      const long id = std::strtol(key.c_str(), &end, 10);
      if (key.empty() || end != key.c_str() + key.size()) throw std::runtime_error("model tree YAML line " + std::to_string(line_no) + ": top-level key '" + key + "' is not a node id");
      cur = &nodes[(int)id];
      if (!val.empty() && val != "{}") throw std::runtime_error("model tree YAML line " + std::to_string(line_no) + ": flow mappings are not supported");
      continue;
    }
    if (!cur) throw std::runtime_error("model tree YAML line " + std::to_string(line_no) + ": field before any node id");
    if (key == "terminal") { cur->terminal = parse_bool(val, line_no); cur->has_terminal = true; }
    else if (key == "model params" || key == "split params") {
      std::vector<double>* dst = key[0] == 'm' ? &cur->model : &cur->split;
      (key[0] == 'm' ? cur->has_model : cur->has_split) = true;
      dst->clear();
      if (val.empty()) seq = dst;
      else if (val.find(']') != std::string::npos) parse_flow_seq(val, *dst, line_no);
      else if (val[0] == '[') { flow_dst = dst; pending_flow = val; flow_line = line_no; }
      else throw std::runtime_error("model tree YAML line " + std::to_string(line_no) + ": expected a sequence after '" + key + "'");
    }
    // other keys ("info", ...) are ignored, like yaml-cpp ignores what the loader never asks for
  }
  if (flow_dst) throw std::runtime_error("model tree YAML: unterminated [ ... ] sequence");
  return nodes;
}

// Linear_model_tree::_build_tree_recursively (model_tree.cc:21-45): ids are handed out by _new_node() in visiting order.
inline void build_flat(const std::map<int, YamlNodeEntry>& y, bool oblique, FlatTree& out, int node_index, int& node_count, int depth = 0) {
  if (depth > 2048) throw std::runtime_error("model tree YAML: tree deeper than 2048 levels");   // bounds the recursion on chain-shaped input
  const int id = out.nodes[node_index].id;
  auto it = y.find(id);
  if (it == y.end() || !it->second.has_terminal) throw std::runtime_error("model tree YAML: node " + std::to_string(id) + " has no 'terminal' field");
  const YamlNodeEntry& e = it->second;
  const std::vector<double>& par = e.terminal ? e.model : e.split;
  if (e.terminal ? !e.has_model : !e.has_split)
    throw std::runtime_error("model tree YAML: node " + std::to_string(id) + " has no '" + (e.terminal ? "model" : "split") + " params'");
  out.nodes[node_index].terminal = e.terminal ? 1 : 0;
  out.nodes[node_index].param_off = (int32_t)out.params.size();
  out.nodes[node_index].n_params = (int32_t)par.size();
  out.params.insert(out.params.end(), par.begin(), par.end());
  if (e.terminal) return;
  if (!oblique && par.size() != 2)   // message of model_tree.cc:33-35
    throw std::runtime_error("Read " + std::to_string(par.size()) + " split parameters for the node " + std::to_string(id) + " instead of the 2 required for a straight tree");
  if (out.nodes.size() > 100000) throw std::runtime_error("model tree YAML: more than 100000 nodes (cyclic ids?)");
  for (int child = 0; child < 2; child++) {
    TreeNode nn; nn.id = ++node_count; nn.terminal = 1; nn.child[0] = nn.child[1] = -1; nn.param_off = 0; nn.n_params = 0;
    out.nodes.push_back(nn);
    const int ci = (int)out.nodes.size() - 1;
    out.nodes[node_index].child[child] = ci;
    build_flat(y, oblique, out, ci, node_count, depth + 1);
  }
}

inline FlatTree parse_model_tree(const char* text, size_t len, bool oblique) {
  const std::map<int, YamlNodeEntry> y = parse_tree_yaml(text, len);
  FlatTree t;
  TreeNode root; root.id = 1; root.terminal = 1; root.child[0] = root.child[1] = -1; root.param_off = 0; root.n_params = 0;
  t.nodes.push_back(root);
  int count = 1;
  build_flat(y, oblique, t, 0, count);
  return t;
}

// number of features a Polynomial_model_tree of this degree builds from n inputs (model_tree.cc:91-123)
inline int poly_feature_count(int n, int degree, bool interaction_only) {
  std::vector<long> cnt(n, 1);   // cnt[i] = number of degree-d monomials whose lowest-index factor is i
  long total = n, prev_total = n;
  std::vector<long> idx(n);
  for (int i = 0; i < n; i++) idx[i] = i;
  for (int d = 1; d < degree; d++) {
    long chunk = 0; std::vector<long> nidx(n, 0);
    for (int i = 0; i < n - (interaction_only ? d : 0); i++) {
      nidx[i] = chunk;
      const long start = idx[i + (interaction_only ? 1 : 0)];
      chunk += prev_total - start;
    }
    for (int i = 0; i < n; i++) idx[i] = nidx[i];
    total += chunk; prev_total = chunk;
    if (total > 100000) return -1;
  }
  return (int)total;
}

// ------------------------------------------------------------------------------------------------ SavedModel variables
// CRC32C (Castagnoli, reflected 0x82F63B78) and TensorFlow's mask (lib/hash/crc32c.h): rot15 + 0xa282ead8.
inline uint32_t crc32c(const unsigned char* p, size_t n) {
  static uint32_t table[256]; static bool ready = false;
  if (!ready) { for (uint32_t i = 0; i < 256; i++) { uint32_t c = i; for (int k = 0; k < 8; k++) c = (c >> 1) ^ (0x82F63B78u & (0u - (c & 1u))); table[i] = c; } ready = true; }
  uint32_t c = 0xFFFFFFFFu;
  for (size_t i = 0; i < n; i++) c = table[(c ^ p[i]) & 0xFF] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}
inline uint32_t crc32c_mask(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }

struct ByteCursor {
  const unsigned char* p; const unsigned char* end;
  uint64_t varint() {
    uint64_t v = 0;
    for (int shift = 0; shift < 64; shift += 7) {
      if (p >= end) throw std::runtime_error("tensor bundle: truncated varint");
      const unsigned char b = *p++;
      v |= (uint64_t)(b & 0x7F) << shift;
      if (!(b & 0x80)) return v;
    }
    throw std::runtime_error("tensor bundle: varint too long");
  }
  const unsigned char* take(size_t n) { if ((size_t)(end - p) < n) throw std::runtime_error("tensor bundle: truncated record"); const unsigned char* q = p; p += n; return q; }
  bool done() const { return p >= end; }
};
inline uint32_t le32(const unsigned char* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

struct BundleEntry { int dtype = 0; std::vector<int64_t> shape; int shard = 0; int64_t offset = 0, size = 0; uint32_t crc = 0; bool has_crc = false, sliced = false; };
struct Bundle { int num_shards = 1; std::map<std::string, BundleEntry> entries; };

// One block of the table (handle = offset,size; the 5-byte trailer = compression type + crc follows it): calls
// visit(key, value bytes) for every entry, undoing the shared-prefix key compression.
template <class F>
inline void table_block(const std::vector<unsigned char>& file, uint64_t off, uint64_t size, F&& visit) {
  if (size < 4 || size > file.size() || file.size() - size < 5 || off > file.size() - size - 5) throw std::runtime_error("tensor bundle index: block handle outside the file");
  if (file[off + size] != 0) throw std::runtime_error("tensor bundle index: compressed block (type " + std::to_string((int)file[off + size]) + "); BundleWriter writes uncompressed tables");
  if (le32(&file[off + size + 1]) != crc32c_mask(crc32c(&file[off], size + 1))) throw std::runtime_error("tensor bundle index: block checksum mismatch");
  const uint32_t n_restarts = le32(&file[off + size - 4]);
  if ((uint64_t)n_restarts * 4 + 4 > size) throw std::runtime_error("tensor bundle index: bad restart array");
  ByteCursor c{&file[off], &file[off] + size - 4 - (size_t)n_restarts * 4};
  std::string key;
  while (!c.done()) {
    const uint64_t shared = c.varint(), non_shared = c.varint(), vlen = c.varint();
    if (shared > key.size()) throw std::runtime_error("tensor bundle index: bad shared-prefix length");
    key.resize(shared);
    const unsigned char* k = c.take(non_shared); key.append((const char*)k, non_shared);
    const unsigned char* v = c.take(vlen);
    visit(key, v, (size_t)vlen);
  }
}

inline BundleEntry parse_bundle_entry(const unsigned char* v, size_t n) {   // BundleEntryProto (tensor_bundle.proto)
  BundleEntry e; ByteCursor c{v, v + n};
  while (!c.done()) {
    const uint64_t tag = c.varint(); const int field = (int)(tag >> 3), wire = (int)(tag & 7);
    if (wire == 0) {
      const uint64_t x = c.varint();
      if (field == 1) e.dtype = (int)x; else if (field == 3) e.shard = (int)x; else if (field == 4) e.offset = (int64_t)x; else if (field == 5) e.size = (int64_t)x;
    } else if (wire == 5) { const unsigned char* q = c.take(4); if (field == 6) { e.crc = le32(q); e.has_crc = true; } }
    else if (wire == 1) c.take(8);
    else if (wire == 2) {
      const uint64_t len = c.varint(); const unsigned char* q = c.take(len);
      if (field == 7) e.sliced = true;
      if (field == 2) {   // TensorShapeProto: repeated Dim dim = 2 { int64 size = 1; string name = 2 }
        ByteCursor s{q, q + len};
        while (!s.done()) {
          const uint64_t t2 = s.varint();
          if ((t2 & 7) == 2) {
            const uint64_t l2 = s.varint(); const unsigned char* d = s.take(l2);
            if ((t2 >> 3) == 2) { ByteCursor dc{d, d + l2}; int64_t dim = 0; while (!dc.done()) { const uint64_t t3 = dc.varint(); if ((t3 & 7) == 0) { const uint64_t x = dc.varint(); if ((t3 >> 3) == 1) dim = (int64_t)x; } else if ((t3 & 7) == 2) dc.take(dc.varint()); else throw std::runtime_error("tensor bundle: bad TensorShapeProto.Dim"); } e.shape.push_back(dim); }
          } else if ((t2 & 7) == 0) s.varint();
          else throw std::runtime_error("tensor bundle: bad TensorShapeProto");
        }
      }
    } else throw std::runtime_error("tensor bundle: unsupported protobuf wire type in BundleEntryProto");
  }
  return e;
}

// <prefix>.index -> every key's entry.  Table layout: data blocks, metaindex block, index block, 48-byte footer
// (metaindex handle, index handle, padding, magic 0xdb4775248b80fb57 little-endian).
inline Bundle read_bundle_index(const std::string& prefix) {
  const std::vector<unsigned char> file = read_file((prefix + ".index").c_str());
  if (file.size() < 48) throw std::runtime_error("tensor bundle index: file shorter than a table footer");
  const unsigned char* foot = &file[file.size() - 48];
  if (le32(foot + 40) != 0x8b80fb57u || le32(foot + 44) != 0xdb477524u) throw std::runtime_error("tensor bundle index: bad table magic (not a TensorFlow checkpoint index)");
  ByteCursor fc{foot, foot + 40};
  fc.varint(); fc.varint();   // metaindex handle
  const uint64_t idx_off = fc.varint(), idx_size = fc.varint();
  Bundle b; bool header_seen = false;
  table_block(file, idx_off, idx_size, [&](const std::string&, const unsigned char* hv, size_t hn) {
    ByteCursor hc{hv, hv + hn}; const uint64_t off = hc.varint(), size = hc.varint();
    table_block(file, off, size, [&](const std::string& key, const unsigned char* v, size_t n) {
      if (key.empty()) {   // BundleHeaderProto: num_shards = 1, endianness = 2 (0 little), version = 3
        header_seen = true; ByteCursor c{v, v + n};
        while (!c.done()) {
          const uint64_t tag = c.varint();
          if ((tag & 7) == 0) { const uint64_t x = c.varint(); if ((tag >> 3) == 1) b.num_shards = (int)x; if ((tag >> 3) == 2 && x != 0) throw std::runtime_error("tensor bundle: big-endian checkpoint"); }
          else if ((tag & 7) == 2) c.take(c.varint());
          else throw std::runtime_error("tensor bundle: bad BundleHeaderProto");
        }
      } else b.entries[key] = parse_bundle_entry(v, n);
    });
  });
  if (!header_seen) throw std::runtime_error("tensor bundle index: no header entry");
  return b;
}

// One DT_FLOAT tensor of the bundle, appended to out; the shape must be exactly `want`.
inline void read_bundle_float(const std::string& prefix, const Bundle& b, const std::string& key, const std::vector<int64_t>& want,
                              std::map<int, std::vector<unsigned char>>& shards, std::vector<float>& out) {
  auto it = b.entries.find(key);
  if (it == b.entries.end()) throw std::runtime_error("SavedModel has no variable '" + key + "'");
  const BundleEntry& e = it->second;
  if (e.dtype != 1) throw std::runtime_error("variable '" + key + "' is not DT_FLOAT (dtype " + std::to_string(e.dtype) + ")");
  if (e.sliced) throw std::runtime_error("variable '" + key + "' is stored as slices");
  if (e.shape != want) {
    std::string got, exp; for (auto d : e.shape) got += std::to_string(d) + ","; for (auto d : want) exp += std::to_string(d) + ",";
    throw std::runtime_error("variable '" + key + "' has shape [" + got + "] instead of [" + exp + "]");
  }
  int64_t count = 1; for (auto d : want) count *= d;
  if (e.size != count * 4) throw std::runtime_error("variable '" + key + "': byte size does not match its shape");
  if (e.shard < 0 || e.shard >= b.num_shards) throw std::runtime_error("variable '" + key + "': shard id out of range");
  if (!shards.count(e.shard)) { char name[64]; std::snprintf(name, sizeof(name), ".data-%05d-of-%05d", e.shard, b.num_shards); shards[e.shard] = read_file((prefix + name).c_str()); }
  const std::vector<unsigned char>& d = shards[e.shard];
  if (e.offset < 0 || (uint64_t)e.offset + (uint64_t)e.size > d.size()) throw std::runtime_error("variable '" + key + "' lies outside its data shard");
  if (e.has_crc && crc32c_mask(crc32c(&d[e.offset], (size_t)e.size)) != e.crc) throw std::runtime_error("variable '" + key + "': CRC32C mismatch");
  const size_t at = out.size(); out.resize(at + (size_t)count);
  std::memcpy(&out[at], &d[e.offset], (size_t)e.size);
}

// The actor of scripts/rover_training_1.py:26-35 out of <dir>/variables/variables.{index,data-*}: Keras names the i-th
// layer that owns weights `layer_with_weights-i`, a Dense layer's variables `kernel` [in,out] and `bias` [out].
inline std::vector<float> read_saved_model_actor(const std::string& dir, const int* dims, int n_layers) {
  const std::string prefix = dir + "/variables/variables";
  const Bundle b = read_bundle_index(prefix);
  std::map<int, std::vector<unsigned char>> shards; std::vector<float> w;
  for (int l = 0; l < n_layers; l++) {
    const std::string base = "layer_with_weights-" + std::to_string(l) + "/";
    read_bundle_float(prefix, b, base + "kernel/.ATTRIBUTES/VARIABLE_VALUE", {dims[l], dims[l + 1]}, shards, w);
    read_bundle_float(prefix, b, base + "bias/.ATTRIBUTES/VARIABLE_VALUE", {dims[l + 1]}, shards, w);
  }
  if (b.entries.count("layer_with_weights-" + std::to_string(n_layers) + "/kernel/.ATTRIBUTES/VARIABLE_VALUE"))
    throw std::runtime_error("SavedModel has more than " + std::to_string(n_layers) + " weight layers: not the actor of rover_training_1.py");
  return w;
}

}  // namespace tfode_fmt
