// grackle-b200: dependency-free reader for the HDF5 dialect used by Grackle's
// data files (there is no libhdf5 on the target image).  Supports superblock v0,
// symbol-table groups, v1 object headers, contiguous/compact un-filtered datasets
// and v1 attributes holding fixed-point or IEEE-float values of either byte order.
// That is exactly what initialize_cloudy_data.c:92-305 and
// initialize_UVbackground_data.c:50-330 of the reference read through libhdf5.
#ifndef GRB200_H5LITE_HPP
#define GRB200_H5LITE_HPP

#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace h5lite {

struct Value {                       // numeric payload converted to native types
  std::vector<uint64_t> dims;
  std::vector<double> f64;           // filled for float and integer types
  std::vector<long long> i64;        // filled for integer types
  bool is_int = false;
  bool ok = false;
  size_t count() const { size_t n = 1; for (auto d : dims) n *= d; return n; }
};

struct Node {
  std::map<std::string, Node> children;
  std::map<std::string, Value> attrs;
  std::vector<std::string> attr_names;   // includes attributes of unsupported type
  Value data;                            // datasets only
  uint64_t storage_size = 0;
  bool is_dataset = false;
  const Node *find(const std::string &path) const;
};

class File {
 public:
  // returns false (and fills err) when the file cannot be read/understood
  bool open(const std::string &path, std::string *err);
  const Node &root() const { return root_; }

 private:
  std::vector<unsigned char> buf_;
  uint64_t base_ = 0;
  Node root_;
  std::string err_;

  template <typename T> T rd(uint64_t off) const;
  bool in_range(uint64_t off, uint64_t n) const { return off + n <= buf_.size() && off + n >= off; }
  struct Dtype { int cls = -1; bool big = false; bool sign = false; uint32_t size = 0; };
  bool parse_dtype(uint64_t off, Dtype *dt);
  bool parse_space(uint64_t off, std::vector<uint64_t> *dims);
  void decode(const Dtype &dt, const std::vector<uint64_t> &dims, uint64_t off, Value *out);
  bool read_object(uint64_t addr, Node *node, int depth);
  bool read_btree(uint64_t node_addr, uint64_t heap_data, Node *node, int depth, int tree_depth);
  unsigned byte(uint64_t off) const;
  bool cstring(uint64_t off, uint64_t max_len, std::string *out) const;
  bool fail(const std::string &m) { if (err_.empty()) err_ = m; return false; }
};

}  // namespace h5lite
#endif
