// grackle-b200: minimal HDF5 reader (see h5lite.hpp for scope).
#include "h5lite.hpp"

#include <cstdio>
#include <cstring>

namespace h5lite {

template <typename T> T File::rd(uint64_t off) const
{
  T v{};
  if (in_range(off, sizeof(T))) std::memcpy(&v, buf_.data() + off, sizeof(T));
  return v;
}

// one byte, 0 beyond the end of the file (a truncated or corrupt file must fail, not fault)
unsigned File::byte(uint64_t off) const { return in_range(off, 1) ? buf_[off] : 0u; }

// NUL-terminated name starting at off, bounded by the end of the file (and by max_len when given)
bool File::cstring(uint64_t off, uint64_t max_len, std::string *out) const
{
  if (!in_range(off, 1)) return false;
  uint64_t room = buf_.size() - off;
  if (max_len && max_len < room) room = max_len;
  const void *nul = std::memchr(buf_.data() + off, 0, room);
  if (!nul && !max_len) return false;                      // runs off the end of the file
  out->assign((const char *)buf_.data() + off, nul ? (size_t)((const unsigned char *)nul - (buf_.data() + off)) : (size_t)room);
  return true;
}

const Node *Node::find(const std::string &path) const
{
  const Node *n = this;
  size_t p = 0;
  while (p < path.size()) {
    while (p < path.size() && path[p] == '/') ++p;
    size_t e = path.find('/', p);
    if (e == std::string::npos) e = path.size();
    if (e == p) break;
    auto it = n->children.find(path.substr(p, e - p));
    if (it == n->children.end()) return nullptr;
    n = &it->second;
    p = e;
  }
  return n;
}

bool File::open(const std::string &path, std::string *err)
{
  std::FILE *f = std::fopen(path.c_str(), "rb");
  if (!f) { if (err) *err = "cannot open " + path; return false; }
  std::fseek(f, 0, SEEK_END);
  long sz = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  buf_.resize(sz > 0 ? sz : 0);
  size_t got = buf_.empty() ? 0 : std::fread(buf_.data(), 1, buf_.size(), f);
  std::fclose(f);
  static const unsigned char magic[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
  bool ok = got == buf_.size() && buf_.size() >= 96 && std::memcmp(buf_.data(), magic, 8) == 0;
  if (ok && buf_[8] != 0) { ok = false; err_ = "unsupported HDF5 superblock version (need v0)"; }
  if (ok && (buf_[13] != 8 || buf_[14] != 8)) { ok = false; err_ = "unsupported offset size"; }
  if (ok) {
    base_ = rd<uint64_t>(24);
    const uint64_t root_header = rd<uint64_t>(56 + 8);   // root symbol-table entry
    ok = read_object(base_ + root_header, &root_, 0);
  } else if (err_.empty()) {
    err_ = "not an HDF5 file: " + path;
  }
  if (!ok && err) *err = err_;
  return ok;
}

bool File::parse_dtype(uint64_t off, Dtype *dt)
{
  if (!in_range(off, 8)) return fail("datatype message out of range");
  const unsigned cv = buf_[off], b0 = buf_[off + 1];
  dt->cls = cv & 0x0F;
  dt->size = rd<uint32_t>(off + 4);
  dt->big = b0 & 1;
  dt->sign = (b0 >> 3) & 1;
  return true;
}

bool File::parse_space(uint64_t off, std::vector<uint64_t> *dims)
{
  if (!in_range(off, 8)) return fail("dataspace message out of range");
  const unsigned ver = buf_[off], rank = buf_[off + 1];
  uint64_t p = off + (ver == 1 ? 8 : 4);
  if (ver != 1 && ver != 2) return fail("unsupported dataspace version");
  dims->clear();
  for (unsigned i = 0; i < rank; ++i) dims->push_back(rd<uint64_t>(p + 8 * i));
  return true;
}

void File::decode(const Dtype &dt, const std::vector<uint64_t> &dims, uint64_t off, Value *out)
{
  out->dims = dims;
  const size_t n = out->count();
  if (!(dt.cls == 0 || dt.cls == 1) || !in_range(off, (uint64_t)n * dt.size)) return;
  if (dt.cls == 1 && dt.size != 8 && dt.size != 4) return;
  if (dt.cls == 0 && dt.size != 8 && dt.size != 4 && dt.size != 2 && dt.size != 1) return;
  out->f64.resize(n);
  if (dt.cls == 0) { out->i64.resize(n); out->is_int = true; }
  for (size_t i = 0; i < n; ++i) {
    unsigned char raw[8] = {0};
    const unsigned char *src = buf_.data() + off + i * dt.size;
    for (uint32_t b = 0; b < dt.size; ++b) raw[b] = dt.big ? src[dt.size - 1 - b] : src[b];
    if (dt.cls == 1) {
      if (dt.size == 8) { double v; std::memcpy(&v, raw, 8); out->f64[i] = v; }
      else { float v; std::memcpy(&v, raw, 4); out->f64[i] = v; }
    } else {
      uint64_t u = 0;
      std::memcpy(&u, raw, 8);
      long long v = (long long)u;
      if (dt.sign && dt.size < 8) {           // sign-extend
        const uint64_t m = 1ull << (8 * dt.size - 1);
        v = (long long)((u ^ m) - m);
      }
      out->i64[i] = v;
      out->f64[i] = (double)v;
    }
  }
  out->ok = true;
}

bool File::read_object(uint64_t addr, Node *node, int depth)
{
  if (depth > 16) return fail("group nesting too deep");
  if (!in_range(addr, 16) || buf_[addr] != 1) return fail("unsupported object header version");
  const unsigned nmsg = rd<uint16_t>(addr + 2);
  const uint32_t hsize = rd<uint32_t>(addr + 8);
  std::vector<std::pair<uint64_t, uint64_t>> blocks{{addr + 16, hsize}};
  Dtype dt;
  std::vector<uint64_t> dims;
  uint64_t layout = 0;
  bool have_dt = false, have_layout = false;
  unsigned seen = 0;
  for (size_t bi = 0; bi < blocks.size() && seen < nmsg; ++bi) {
    uint64_t p = blocks[bi].first;
    const uint64_t end = p + blocks[bi].second;
    while (p + 8 <= end && seen < nmsg) {
      if (!in_range(p, 8)) return fail("object header runs past end of file");
      const unsigned mtype = rd<uint16_t>(p), msize = rd<uint16_t>(p + 2);
      const uint64_t body = p + 8;
      ++seen;
      switch (mtype) {
        case 0x01: if (!parse_space(body, &dims)) return false; break;
        case 0x03: if (!parse_dtype(body, &dt)) return false; have_dt = true; break;
        case 0x08: layout = body; have_layout = true; break;
        case 0x10: blocks.push_back({base_ + rd<uint64_t>(body), rd<uint64_t>(body + 8)}); break;
        case 0x11: {
          const uint64_t btree = base_ + rd<uint64_t>(body), heap = base_ + rd<uint64_t>(body + 8);
          if (!in_range(heap, 32) || std::memcmp(buf_.data() + heap, "HEAP", 4) != 0)
            return fail("bad local heap");
          if (!read_btree(btree, base_ + rd<uint64_t>(heap + 24), node, depth, 0)) return false;
          break;
        }
        case 0x0C: {
          if (!in_range(body, 8)) return fail("attribute message out of range");
          if (buf_[body] != 1) break;
          auto pad8 = [](uint64_t x) { return (x + 7) & ~7ull; };
          const unsigned nsz = rd<uint16_t>(body + 2), dtsz = rd<uint16_t>(body + 4),
                         dssz = rd<uint16_t>(body + 6);
          uint64_t q = body + 8;
          std::string name;
          if (nsz == 0 || !in_range(q, nsz) || !cstring(q, nsz, &name)) return fail("attribute name out of range");
          q += pad8(nsz);
          Dtype adt;
          std::vector<uint64_t> adims;
          if (!parse_dtype(q, &adt)) return false;
          q += pad8(dtsz);
          if (!parse_space(q, &adims)) return false;
          q += pad8(dssz);
          node->attr_names.push_back(name);
          Value v;
          decode(adt, adims, q, &v);
          if (v.ok) node->attrs[name] = v;
          break;
        }
        default: break;
      }
      p = body + msize;
    }
  }
  if (have_dt && have_layout) {
    node->is_dataset = true;
    node->data.dims = dims;
    const size_t n = node->data.count();
    if (!in_range(layout, 18)) return fail("layout message out of range");
    const unsigned ver = buf_[layout];
    uint64_t data_off = ~0ull;
    if (ver == 3) {
      const unsigned cls = buf_[layout + 1];
      if (cls == 1) { data_off = rd<uint64_t>(layout + 2); node->storage_size = rd<uint64_t>(layout + 10);
                      if (data_off != ~0ull) data_off += base_; }
      else if (cls == 0) { data_off = layout + 4; node->storage_size = rd<uint16_t>(layout + 2); }
      else return fail("chunked datasets are not supported");
    } else if (ver == 1 || ver == 2) {
      if (buf_[layout + 2] != 1) return fail("unsupported dataset layout");
      data_off = base_ + rd<uint64_t>(layout + 8);
      node->storage_size = (uint64_t)n * dt.size;
    } else {
      return fail("unsupported layout message version");
    }
    if (data_off != ~0ull) decode(dt, dims, data_off, &node->data);
  }
  return true;
}

bool File::read_btree(uint64_t na, uint64_t heap_data, Node *node, int depth, int tree_depth)
{
  if (tree_depth > 8) return fail("B-tree too deep (corrupt file?)");
  if (!in_range(na, 24) || std::memcmp(buf_.data() + na, "TREE", 4) != 0) return fail("bad B-tree node");
  const unsigned level = buf_[na + 5], nent = rd<uint16_t>(na + 6);
  if (!in_range(na + 24, 8 + 16ull * nent)) return fail("B-tree node runs past end of file");
  for (unsigned i = 0; i < nent; ++i) {
    const uint64_t child = base_ + rd<uint64_t>(na + 24 + 8 + 16ull * i);
    if (level > 0) {
      if (!read_btree(child, heap_data, node, depth, tree_depth + 1)) return false;
      continue;
    }
    if (!in_range(child, 8) || std::memcmp(buf_.data() + child, "SNOD", 4) != 0)
      return fail("bad symbol-table node");
    const unsigned nsym = rd<uint16_t>(child + 6);
    if (!in_range(child + 8, 40ull * nsym)) return fail("symbol-table node runs past end of file");
    for (unsigned s = 0; s < nsym; ++s) {
      const uint64_t e = child + 8 + 40ull * s;
      const uint64_t name_off = heap_data + rd<uint64_t>(e);
      std::string name;
      if (!cstring(name_off, 0, &name)) return fail("symbol name out of range");
      if (!read_object(base_ + rd<uint64_t>(e + 8), &node->children[name], depth + 1)) return false;
    }
  }
  return true;
}

}  // namespace h5lite
