// =====================================================================================
// nsdf_h5.hpp -- header-only reader for the HDF5 subset Keras `save_weights` files use.
//
// Replaces libhdf5 (an un-vendored system library the reference reaches through
// HighFive 2.1, src/neuralNetwork.cpp:85-151, src/simpleInfer.cpp:13-79) for the one job
// the hot path needs: walking /<layer>/<layer>/{bias:0,kernel:0} and reading float32 data.
// Written from the public "HDF5 File Format Specification Version 2.0"; nothing here is
// derived from libhdf5 or HighFive sources.
//
// Understands: superblock v0/v1, 8-byte (or 4-byte) offsets and lengths, v1 object headers
// with continuation blocks, old-style groups (symbol-table message -> v1 B-tree "TREE" ->
// "SNOD" symbol nodes -> local "HEAP" for names), dataspace v1/v2, IEEE little-endian
// float32 / float64 datatypes, layout v3 contiguous and compact, attribute messages v1-v3 holding
// fixed-length string arrays (Keras' `layer_names`, `weight_names`).  Anything else throws
// nsdf::h5::FormatError (attributes of other types are skipped).
//
// write_keras_weights() emits the same subset (superblock v0, old-style groups, contiguous float32
// datasets, `layer_names` / `weight_names` / `backend` attributes): the synthetic models of the scaling and
// parity tests (other depths, 4-input animation nets) are real files that go through the same reader.
// =====================================================================================
#ifndef NSDF_H5_HPP_
#define NSDF_H5_HPP_

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "nsdf_inflate.hpp"

namespace nsdf {
namespace h5 {

struct FormatError : std::runtime_error {
    explicit FormatError(const std::string& what) : std::runtime_error("nsdf::h5: " + what) {}
};

enum class NodeKind { Group, Dataset };

struct Node {
    NodeKind kind = NodeKind::Group;
    std::string name;
    // group: children keyed by link name; std::map iterates in byte-wise ascending name order,
    // which is the order HighFive::listObjectNames returns (H5_INDEX_NAME, H5_ITER_INC).
    std::map<std::string, std::shared_ptr<Node>> children;
    // dataset
    std::vector<size_t> dims;
    std::vector<float> data;      // converted to float32
    int64_t file_offset = -1;     // absolute offset of contiguous raw data, -1 if compact
    // attributes that are arrays of fixed-length strings (layer_names, weight_names), NUL padding stripped
    std::map<std::string, std::vector<std::string>> string_attrs;
};

class File {
public:
    explicit File(const std::string& path) {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) throw FormatError("unable to open file '" + path + "'");
        std::fseek(f, 0, SEEK_END);
        const long n = std::ftell(f);
        std::fseek(f, 0, SEEK_SET);
        buf_.resize(n > 0 ? (size_t)n : 0);
        const size_t got = buf_.empty() ? 0 : std::fread(buf_.data(), 1, buf_.size(), f);
        std::fclose(f);
        if (got != buf_.size()) throw FormatError("short read on '" + path + "'");
        parse_superblock();
    }
    File(const uint8_t* bytes, size_t n) : buf_(bytes, bytes + n) { parse_superblock(); }

    const Node& root() const { return *root_; }
    std::shared_ptr<Node> root_ptr() const { return root_; }

private:
    std::vector<uint8_t> buf_;
    std::shared_ptr<Node> root_;
    unsigned so_ = 8, sl_ = 8;   // size of offsets / lengths
    uint64_t base_ = 0;
    int depth_ = 0;

    void need(uint64_t off, uint64_t n) const {
        if (off > buf_.size() || n > buf_.size() - off) throw FormatError("structure runs past end of file");
    }
    uint64_t rd(uint64_t off, unsigned n) const {
        need(off, n);
        uint64_t v = 0;
        for (unsigned i = 0; i < n; ++i) v |= (uint64_t)buf_[off + i] << (8 * i);
        return v;
    }
    bool sig(uint64_t off, const char* s) const {
        need(off, 4);
        return std::memcmp(buf_.data() + off, s, 4) == 0;
    }

    void parse_superblock() {
        static const uint8_t magic[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
        if (buf_.size() < 96 || std::memcmp(buf_.data(), magic, 8) != 0) throw FormatError("not an HDF5 file");
        const unsigned ver = buf_[8];
        if (ver > 1) throw FormatError("unsupported superblock version " + std::to_string(ver));
        so_ = buf_[13];
        sl_ = buf_[14];
        if ((so_ != 8 && so_ != 4) || (sl_ != 8 && sl_ != 4)) throw FormatError("unsupported offset/length size");
        uint64_t p = 24 + (ver == 1 ? 4 : 0);
        base_ = rd(p, so_);
        p += 4 * so_;   // base, free-space, end-of-file, driver-info addresses
        // root group symbol-table entry: link-name offset, object-header address, cache type, ...
        const uint64_t root_hdr = rd(p + so_, so_) + base_;
        root_ = read_object("/", root_hdr);
        if (root_->kind != NodeKind::Group) throw FormatError("root object is not a group");
    }

    struct Msg { unsigned type; uint64_t body; unsigned size; };

    std::vector<Msg> messages(uint64_t addr) const {
        if (rd(addr, 1) != 1) throw FormatError("unsupported object-header version");
        const unsigned nmsg = (unsigned)rd(addr + 2, 2);
        const uint64_t hsize = rd(addr + 8, 4);
        std::vector<std::pair<uint64_t, uint64_t>> blocks;
        blocks.emplace_back(addr + 16, hsize);   // 12-byte prefix padded to 8
        std::vector<Msg> out;
        for (size_t bi = 0; bi < blocks.size() && out.size() < nmsg; ++bi) {
            uint64_t pos = blocks[bi].first;
            const uint64_t end = pos + blocks[bi].second;
            need(pos, blocks[bi].second);
            while (pos + 8 <= end && out.size() < nmsg) {
                Msg m;
                m.type = (unsigned)rd(pos, 2);
                m.size = (unsigned)rd(pos + 2, 2);
                m.body = pos + 8;
                need(m.body, m.size);
                if (m.type == 0x0010) blocks.emplace_back(rd(m.body, so_) + base_, rd(m.body + so_, sl_));
                out.push_back(m);
                pos = m.body + m.size;
            }
            if (blocks.size() > 4096) throw FormatError("object-header continuation loop");
        }
        return out;
    }

    std::string heap_name(uint64_t heap, uint64_t off) const {
        if (!sig(heap, "HEAP")) throw FormatError("bad local-heap signature");
        const uint64_t data = rd(heap + 8 + 2 * sl_, so_) + base_;
        uint64_t s = data + off;
        need(s, 1);
        const void* e = std::memchr(buf_.data() + s, 0, buf_.size() - s);
        if (!e) throw FormatError("unterminated link name");
        return std::string((const char*)buf_.data() + s, (const char*)e);
    }

    void walk_btree(uint64_t addr, uint64_t heap, std::vector<std::pair<std::string, uint64_t>>& out, int level) const {
        if (level > 32) throw FormatError("B-tree too deep");
        if (sig(addr, "TREE")) {
            if (rd(addr + 4, 1) != 0) throw FormatError("expected a group B-tree node");
            const unsigned used = (unsigned)rd(addr + 6, 2);
            uint64_t pos = addr + 8 + 2 * so_;   // skip left/right sibling
            for (unsigned i = 0; i < used; ++i) {
                const uint64_t child = rd(pos + sl_, so_) + base_;   // key_i (sl_) then child_i (so_)
                walk_btree(child, heap, out, level + 1);
                pos += sl_ + so_;
            }
        } else if (sig(addr, "SNOD")) {
            const unsigned nsym = (unsigned)rd(addr + 6, 2);
            uint64_t pos = addr + 8;
            const uint64_t entry = 2 * so_ + 4 + 4 + 16;
            for (unsigned i = 0; i < nsym; ++i) {
                out.emplace_back(heap_name(heap, rd(pos, so_)), rd(pos + so_, so_) + base_);
                pos += entry;
            }
        } else {
            throw FormatError("unexpected node signature in group B-tree");
        }
    }

    std::shared_ptr<Node> read_object(const std::string& name, uint64_t hdr) {
        if (++depth_ > 64) throw FormatError("group nesting too deep");
        auto node = std::make_shared<Node>();
        node->name = name;
        const std::vector<Msg> msgs = messages(hdr);
        const Msg* stab = nullptr;
        const Msg* layout = nullptr;
        for (const Msg& m : msgs) {
            if (m.type == 0x0011) stab = &m;
            if (m.type == 0x0008) layout = &m;
            if (m.type == 0x000C) read_string_attr(*node, m);
        }
        if (stab) {
            node->kind = NodeKind::Group;
            std::vector<std::pair<std::string, uint64_t>> entries;
            walk_btree(rd(stab->body, so_) + base_, rd(stab->body + so_, so_) + base_, entries, 0);
            for (auto& e : entries) node->children[e.first] = read_object(e.first, e.second);
        } else if (layout) {
            node->kind = NodeKind::Dataset;
            read_dataset(*node, msgs);
        } else {
            throw FormatError("object '" + name + "' is neither an old-style group nor a dataset");
        }
        --depth_;
        return node;
    }

    // Attribute message (0x000C) versions 1-3: keeps arrays (or scalars) of fixed-length strings, ignores everything else.
    void read_string_attr(Node& node, const Msg& m) const {
        if (m.size < 8) return;
        const unsigned ver = (unsigned)rd(m.body, 1);
        if (ver < 1 || ver > 3) return;
        const unsigned name_sz = (unsigned)rd(m.body + 2, 2), dt_sz = (unsigned)rd(m.body + 4, 2), ds_sz = (unsigned)rd(m.body + 6, 2);
        const auto pad = [&](uint64_t v) { return ver == 1 ? (v + 7u) & ~(uint64_t)7u : v; };
        uint64_t p = m.body + 8 + (ver == 3 ? 1 : 0);
        const uint64_t end = m.body + m.size;
        if (name_sz == 0 || p + pad(name_sz) + pad(dt_sz) + pad(ds_sz) > end) return;
        const std::string name((const char*)buf_.data() + p, strnlen((const char*)buf_.data() + p, name_sz));
        p += pad(name_sz);
        const uint64_t dt = p;
        p += pad(dt_sz);
        const uint64_t ds = p;
        p += pad(ds_sz);
        if (dt_sz < 8 || ((unsigned)rd(dt, 1) & 0x0F) != 3) return;   // class 3 = fixed-length string
        const uint64_t len = rd(dt + 4, 4);
        if (len == 0 || ds_sz < 4) return;
        const unsigned dsver = (unsigned)rd(ds, 1), rank = (unsigned)rd(ds + 1, 1);
        if ((dsver != 1 && dsver != 2) || rank > 1) return;
        uint64_t count = 1;
        if (rank == 1) {
            const uint64_t dims = ds + (dsver == 1 ? 8 : 4);
            if (dims + sl_ > ds + ds_sz) return;
            count = rd(dims, sl_);
        }
        if (count > (end - p) / len) return;
        std::vector<std::string> v;
        for (uint64_t i = 0; i < count; ++i) {
            const char* c = (const char*)buf_.data() + p + i * len;
            v.emplace_back(c, strnlen(c, (size_t)len));
        }
        node.string_attrs[name] = std::move(v);
    }

    // Raw-data chunk B-tree (version 1, node type 1): collects (chunk offsets in elements, file address, stored size, filter mask)
    struct Chunk { std::vector<uint64_t> off; uint64_t addr; uint32_t size; uint32_t mask; };
    void walk_chunk_btree(uint64_t addr, unsigned nd, std::vector<Chunk>& out, int depth) const {
        if (depth > 16) throw FormatError("chunk B-tree too deep");
        need(addr, 8 + 2 * so_);
        if (!sig(addr, "TREE") || rd(addr + 4, 1) != 1) throw FormatError("bad chunk B-tree node");
        const unsigned level = (unsigned)rd(addr + 5, 1), used = (unsigned)rd(addr + 6, 2);
        const uint64_t key = 8 + 8ull * nd;                 // chunk size (4), filter mask (4), nd offsets of 8 bytes
        uint64_t pos = addr + 8 + 2 * so_;
        for (unsigned i = 0; i < used; ++i) {
            need(pos, key + so_);
            const uint64_t child = rd(pos + key, so_) + base_;
            if (level > 0) {
                walk_chunk_btree(child, nd, out, depth + 1);
            } else {
                Chunk c;
                c.size = (uint32_t)rd(pos, 4);
                c.mask = (uint32_t)rd(pos + 4, 4);
                for (unsigned k = 0; k < nd; ++k) c.off.push_back(rd(pos + 8 + 8ull * k, 8));
                c.addr = child;
                out.push_back(std::move(c));
            }
            pos += key + so_;
        }
    }
    void read_chunked(Node& d, uint64_t n, unsigned elem, uint64_t btree, const std::vector<uint64_t>& cdims,
                      const std::vector<unsigned>& filters) const {
        const size_t rank = d.dims.size();
        if (cdims.size() != rank + 1 || cdims[rank] != elem) throw FormatError("dataset '" + d.name + "': bad chunk dimensions");
        uint64_t chunk_elems = 1;
        for (size_t k = 0; k < rank; ++k) {
            if (cdims[k] == 0 || chunk_elems > (1ull << 28) / cdims[k]) throw FormatError("dataset '" + d.name + "': chunk too large");
            chunk_elems *= cdims[k];
        }
        for (unsigned f : filters)
            if (f != 1 && f != 2 && f != 3) throw FormatError("dataset '" + d.name + "': unsupported filter " + std::to_string(f));
        d.data.assign((size_t)n, 0.0f);           // unallocated chunks read as the fill value 0
        if (btree == ~(uint64_t)0 + base_ || n == 0) return;
        std::vector<Chunk> chunks;
        walk_chunk_btree(btree, (unsigned)rank + 1, chunks, 0);
        std::vector<uint8_t> raw;
        for (const Chunk& c : chunks) {
            need(c.addr, c.size);
            raw.assign(buf_.begin() + (size_t)c.addr, buf_.begin() + (size_t)(c.addr + c.size));
            for (size_t fi = filters.size(); fi-- > 0;) {          // undo the pipeline back to front
                if (c.mask & (1u << fi)) continue;                 // this filter was skipped for this chunk
                if (filters[fi] == 1) {
                    raw = zl::zlib_inflate(raw.data(), raw.size(), (size_t)chunk_elems * elem);
                } else if (filters[fi] == 2) {                     // shuffle: byte b of element e was stored at b * count + e
                    const size_t cnt = raw.size() / elem;
                    std::vector<uint8_t> un(raw.size());
                    for (size_t e = 0; e < cnt; ++e)
                        for (unsigned b = 0; b < elem; ++b) un[e * elem + b] = raw[(size_t)b * cnt + e];
                    for (size_t t = cnt * elem; t < raw.size(); ++t) un[t] = raw[t];
                    raw.swap(un);
                } else if (filters[fi] == 3) {                     // fletcher32: four checksum bytes behind the data
                    if (raw.size() < 4) throw FormatError("dataset '" + d.name + "': short fletcher32 chunk");
                    raw.resize(raw.size() - 4);
                }
            }
            if (raw.size() < chunk_elems * elem) throw FormatError("dataset '" + d.name + "': short chunk");
            // scatter the chunk into the row-major dataset, clipping at the extent
            std::vector<uint64_t> idx(rank, 0);
            for (uint64_t e = 0; e < chunk_elems; ++e) {
                bool inside = true;
                uint64_t lin = 0;
                for (size_t k = 0; k < rank; ++k) {
                    const uint64_t g = c.off[k] + idx[k];
                    if (g >= d.dims[k]) { inside = false; break; }
                    lin = lin * d.dims[k] + g;
                }
                if (inside) {
                    if (elem == 4) {
                        std::memcpy(&d.data[(size_t)lin], raw.data() + e * 4, 4);
                    } else {
                        double v;
                        std::memcpy(&v, raw.data() + e * 8, 8);
                        d.data[(size_t)lin] = (float)v;
                    }
                }
                for (size_t k = rank; k-- > 0;) {
                    if (++idx[k] < cdims[k]) break;
                    idx[k] = 0;
                }
            }
        }
    }

    void read_dataset(Node& d, const std::vector<Msg>& msgs) const {
        bool have_dims = false, have_layout = false;
        unsigned elem = 0;   // 4 = float32, 8 = float64
        uint64_t data_off = 0, data_size = 0;
        bool compact = false, chunked = false;
        uint64_t chunk_btree = 0;
        std::vector<uint64_t> chunk_dims;        // rank + 1 entries, the last one is the element size
        std::vector<unsigned> filters;           // filter ids in pipeline order (1 = deflate, 2 = shuffle, 3 = fletcher32)
        for (const Msg& m : msgs) {
            if (m.type == 0x0001) {
                const unsigned ver = (unsigned)rd(m.body, 1), rank = (unsigned)rd(m.body + 1, 1);
                uint64_t p;
                if (ver == 1) p = m.body + 8;
                else if (ver == 2) p = m.body + 4;
                else throw FormatError("unsupported dataspace version");
                d.dims.clear();
                for (unsigned i = 0; i < rank; ++i) d.dims.push_back((size_t)rd(p + (uint64_t)i * sl_, sl_));
                have_dims = true;
            } else if (m.type == 0x0003) {
                const unsigned cls = (unsigned)rd(m.body, 1) & 0x0F;
                const unsigned bits0 = (unsigned)rd(m.body + 1, 1);
                const unsigned size = (unsigned)rd(m.body + 4, 4);
                if (cls == 1 && (bits0 & 1) == 0) {
                    const unsigned prec = (unsigned)rd(m.body + 10, 2), epos = (unsigned)rd(m.body + 12, 1),
                                   esize = (unsigned)rd(m.body + 13, 1), mpos = (unsigned)rd(m.body + 14, 1),
                                   msz = (unsigned)rd(m.body + 15, 1);
                    const uint64_t ebias = rd(m.body + 16, 4);
                    if (size == 4 && prec == 32 && epos == 23 && esize == 8 && mpos == 0 && msz == 23 && ebias == 127)
                        elem = 4;
                    if (size == 8 && prec == 64 && epos == 52 && esize == 11 && mpos == 0 && msz == 52 && ebias == 1023)
                        elem = 8;
                }
            } else if (m.type == 0x0008) {
                if (rd(m.body, 1) != 3) throw FormatError("unsupported data-layout message version");
                const unsigned lclass = (unsigned)rd(m.body + 1, 1);
                if (lclass == 1) {
                    data_off = rd(m.body + 2, so_) + base_;
                    data_size = rd(m.body + 2 + so_, sl_);
                } else if (lclass == 0) {
                    data_size = rd(m.body + 2, 2);
                    data_off = m.body + 4;
                    compact = true;
                } else if (lclass == 2) {
                    // chunked storage (what h5py writes for compressed or resizable datasets): rank + 1 chunk dimensions
                    // (the last one is the element size) and a version-1 B-tree of raw chunks
                    const unsigned cr = (unsigned)rd(m.body + 2, 1);
                    chunk_btree = rd(m.body + 3, so_) + base_;
                    for (unsigned i = 0; i < cr; ++i) chunk_dims.push_back(rd(m.body + 3 + so_ + 4u * i, 4));
                    chunked = true;
                } else {
                    throw FormatError("dataset '" + d.name + "': unknown storage layout class");
                }
                have_layout = true;
            } else if (m.type == 0x000B) {
                // filter pipeline, version 1 or 2
                const unsigned ver = (unsigned)rd(m.body, 1), nf = (unsigned)rd(m.body + 1, 1);
                uint64_t p = m.body + (ver == 1 ? 8 : 2);
                for (unsigned i = 0; i < nf; ++i) {
                    const unsigned id = (unsigned)rd(p, 2);
                    unsigned name_len = 0;
                    if (ver == 1 || id >= 256) { name_len = (unsigned)rd(p + 2, 2); p += 2; }
                    const unsigned ncv = (unsigned)rd(p + 4, 2);
                    p += 6;
                    if (ver == 1) name_len = (name_len + 7u) & ~7u;
                    p += name_len + 4u * ncv;
                    if (ver == 1 && (ncv & 1u)) p += 4;
                    filters.push_back(id);
                }
            }
        }
        if (!have_dims || !have_layout) throw FormatError("dataset '" + d.name + "' lacks dataspace or layout");
        if (!elem) throw FormatError("dataset '" + d.name + "': only little-endian IEEE float32/float64 supported");
        uint64_t n = 1;
        for (size_t v : d.dims) {
            if (v != 0 && n > (uint64_t)buf_.size() / v) throw FormatError("dataset '" + d.name + "' larger than file");
            n *= v;
        }
        if (chunked) {
            read_chunked(d, n, elem, chunk_btree, chunk_dims, filters);
            return;
        }
        if (data_size < n * elem) throw FormatError("dataset '" + d.name + "': storage smaller than extent");
        need(data_off, n * elem);
        d.data.resize((size_t)n);
        if (elem == 4) {
            if (n) std::memcpy(d.data.data(), buf_.data() + data_off, (size_t)n * 4);
        } else {
            for (uint64_t i = 0; i < n; ++i) {
                double v;
                std::memcpy(&v, buf_.data() + data_off + 8 * i, 8);
                d.data[(size_t)i] = (float)v;
            }
        }
        d.file_offset = compact ? -1 : (int64_t)data_off;
    }
};

// One dense layer as NeuralNetwork::load sees it (src/neuralNetwork.cpp:99-131).
struct KerasLayer {
    std::string name;
    std::vector<size_t> kernel_dims;   // [in][out]
    std::vector<float> kernel;         // row-major [in][out]
    std::vector<float> bias;           // [out]
    int64_t kernel_offset = -1, bias_offset = -1;
};

// Walks the file like NeuralNetwork::load: top-level names ascending; each must be a group with
// exactly one child group; inside, rank-1 datasets are biases, rank-2 kernels.  On a structure the
// reference rejects, `error` receives its message ("Unsupported Layer", ...) and false is returned.
//
// Order of the layers: the reference takes the name-sorted order (listObjectNames), which is the network's order only up to
// ten layers ("dense_10" sorts before "dense_2").  When the root carries Keras' `layer_names` attribute and it names exactly
// the groups present, that order is used instead -- identical for every file the reference can load correctly, right for
// deeper ones.
inline std::vector<std::string> keras_layer_order(const Node& root) {
    std::vector<std::string> sorted;
    for (const auto& kv : root.children) sorted.push_back(kv.first);
    const auto it = root.string_attrs.find("layer_names");
    if (it == root.string_attrs.end() || it->second.size() != sorted.size()) return sorted;
    std::vector<std::string> named = it->second;
    std::sort(named.begin(), named.end());
    if (named != sorted) return sorted;   // (sorted is ascending already; duplicates or strangers fail the comparison)
    return it->second;
}

inline bool read_keras_layers(const File& f, std::vector<KerasLayer>& out, std::string& error) {
    out.clear();
    for (const std::string& lname : keras_layer_order(f.root())) {
        const auto kv = *f.root().children.find(lname);
        const Node& g = *kv.second;
        if (g.kind != NodeKind::Group) { error = "Unsupported Layer"; return false; }
        if (g.children.size() != 1) { error = "Unsupported Layer"; return false; }
        auto it = g.children.find(kv.first);
        if (it == g.children.end() || it->second->kind != NodeKind::Group) { error = "Unsupported Layer"; return false; }
        KerasLayer L;
        L.name = kv.first;
        for (const auto& dv : it->second->children) {
            const Node& d = *dv.second;
            if (d.kind != NodeKind::Dataset) { error = "Unsupported Layer"; return false; }
            if (d.dims.size() == 1) { L.bias = d.data; L.bias_offset = d.file_offset; }
            else if (d.dims.size() == 2) { L.kernel = d.data; L.kernel_dims = d.dims; L.kernel_offset = d.file_offset; }
            else { error = "Unsupported layer, to many dims!"; return false; }
        }
        if (L.kernel_dims.size() != 2 || L.bias.size() != L.kernel_dims[1]) { error = "Unsupported Layer"; return false; }
        out.push_back(std::move(L));
    }
    if (out.empty()) { error = "no layers in file"; return false; }
    return true;
}


// ---------------------------------------------------------------------------------------------
// Writer for the same subset: what Keras' save_weights + libhdf5 1.10 lay down for a stack of
// Dense layers, structure by structure (superblock v0, v1 object headers, symbol-table groups with a level-0
// "TREE", "SNOD" leaves of at most 2*4 entries, local "HEAP" with a trailing free block, contiguous
// little-endian float32 datasets with dataspace v1 / datatype v1 / fill-value v2 / layout v3 messages,
// fixed-length-string attributes `layer_names` and `weight_names`).  Byte patterns follow the HDF5 File
// Format Specification 2.0 and were compared field by field with the shipped data/neuralGeometries files.
class Writer {
public:
    // kernel: [in][out] row-major (Keras); bias: [out].  Layers are written in the given order; the group of
    // layer i is names[i] (e.g. dense, dense_1, ...).
    // storage: 0 = contiguous (what Keras / h5py write for these small datasets, and what the shipped files hold),
    //          1 = chunked, 2 = chunked + shuffle + deflate (zlib streams of stored blocks: the layout h5py produces
    //          with compression="gzip", shuffle=True, without a compressor of our own)
    static std::vector<uint8_t> keras_weights(const std::vector<KerasLayer>& layers, int storage = 0) {
        if (layers.empty()) throw FormatError("no layers to write");
        Writer w;
        w.storage_ = storage;
        w.buf_.assign(96, 0);
        std::vector<std::pair<std::string, Child>> top;
        std::vector<std::string> layer_names;
        for (const KerasLayer& L : layers) {
            if (L.kernel_dims.size() != 2 || L.kernel.size() != L.kernel_dims[0] * L.kernel_dims[1] ||
                L.bias.size() != L.kernel_dims[1] || L.name.empty())
                throw FormatError("layer '" + L.name + "': kernel [in][out] and bias [out] do not agree");
            const Child bias = w.dataset({L.bias.size()}, L.bias);
            const Child kernel = w.dataset({L.kernel_dims[0], L.kernel_dims[1]}, L.kernel);
            const Child inner = w.group({{"bias:0", bias}, {"kernel:0", kernel}}, {});
            const Child outer = w.group({{L.name, inner}},
                                        {string_attr("weight_names", {L.name + "/kernel:0", L.name + "/bias:0"})});
            top.emplace_back(L.name, outer);
            layer_names.push_back(L.name);
        }
        for (size_t i = 0; i < layer_names.size(); ++i)
            for (size_t j = i + 1; j < layer_names.size(); ++j)
                if (layer_names[i] == layer_names[j]) throw FormatError("duplicate layer name '" + layer_names[i] + "'");
        std::sort(top.begin(), top.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
        const Child root = w.group(top, {string_attr("layer_names", layer_names), string_attr("backend", {"tensorflow"})});
        // superblock v0
        static const uint8_t magic[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
        std::memcpy(w.buf_.data(), magic, 8);
        w.buf_[13] = 8; w.buf_[14] = 8;            // sizes of offsets and lengths
        w.put(16, LEAF_K, 2); w.put(18, INTERNAL_K, 2);
        w.put(24, 0, 8);                            // base address
        w.put(32, UNDEF, 8);                        // free-space info
        w.put(40, w.buf_.size(), 8);                // end of file
        w.put(48, UNDEF, 8);                        // driver info
        w.put(56, 0, 8); w.put(64, root.header, 8); // root symbol-table entry: name offset, object header
        w.put(72, 1, 4);                            // cache type 1: scratch holds B-tree and heap addresses
        w.put(80, root.btree, 8); w.put(88, root.heap, 8);
        return std::move(w.buf_);
    }

    static void keras_weights(const std::string& path, const std::vector<KerasLayer>& layers, int storage = 0) {
        const std::vector<uint8_t> bytes = keras_weights(layers, storage);
        FILE* f = std::fopen(path.c_str(), "wb");
        if (!f) throw FormatError("unable to create file '" + path + "'");
        const size_t put = std::fwrite(bytes.data(), 1, bytes.size(), f);
        if (std::fclose(f) != 0 || put != bytes.size()) throw FormatError("short write on '" + path + "'");
    }

private:
    static constexpr uint64_t UNDEF = ~(uint64_t)0;
    static constexpr unsigned LEAF_K = 4, INTERNAL_K = 16;
    struct Child { uint64_t header = 0, btree = 0, heap = 0; bool is_group = false; };
    typedef std::vector<uint8_t> Bytes;   // one header message: type(2) size(2) flags(1) reserved(3) body (multiple of 8)
    std::vector<uint8_t> buf_;
    int storage_ = 0;

    uint64_t alloc(size_t n) {
        const uint64_t off = (buf_.size() + 7u) & ~(uint64_t)7u;
        buf_.resize(off + n, 0);
        return off;
    }
    void put(uint64_t off, uint64_t v, unsigned n) {
        for (unsigned i = 0; i < n; ++i) buf_[off + i] = (uint8_t)(v >> (8 * i));
    }
    static void app(Bytes& b, uint64_t v, unsigned n) {
        for (unsigned i = 0; i < n; ++i) b.push_back((uint8_t)(v >> (8 * i)));
    }
    static void pad8(Bytes& b) { while (b.size() & 7u) b.push_back(0); }
    static Bytes message(unsigned type, unsigned flags, Bytes body) {
        pad8(body);
        Bytes m;
        app(m, type, 2); app(m, body.size(), 2); app(m, flags, 1); app(m, 0, 3);
        m.insert(m.end(), body.begin(), body.end());
        return m;
    }
    static Bytes string_attr(const std::string& name, const std::vector<std::string>& values) {
        size_t len = 1;
        for (const std::string& v : values) len = std::max(len, v.size());
        Bytes b;
        app(b, 1, 1); app(b, 0, 1); app(b, name.size() + 1, 2); app(b, 8, 2); app(b, 24, 2);
        b.insert(b.end(), name.begin(), name.end()); b.push_back(0); pad8(b);
        app(b, 0x13, 1); app(b, 0x01, 1); app(b, 0, 2); app(b, len, 4);                  // string, NUL-padded, ASCII
        app(b, 1, 1); app(b, 1, 1); app(b, 1, 1); app(b, 0, 5); app(b, values.size(), 8); app(b, values.size(), 8);
        for (const std::string& v : values) {
            b.insert(b.end(), v.begin(), v.end());
            b.insert(b.end(), len - v.size(), 0);
        }
        return message(0x000C, 4, b);
    }
    uint64_t object_header(const std::vector<Bytes>& msgs) {
        size_t total = 0;
        for (const Bytes& m : msgs) total += m.size();
        const uint64_t at = alloc(16 + total);
        put(at, 1, 1); put(at + 2, msgs.size(), 2); put(at + 4, 1, 4); put(at + 8, total, 4);
        uint64_t p = at + 16;
        for (const Bytes& m : msgs) { std::memcpy(buf_.data() + p, m.data(), m.size()); p += m.size(); }
        return at;
    }
    static uint32_t adler32(const uint8_t* p, size_t n) {
        uint32_t a = 1, b = 0;
        for (size_t i = 0; i < n; ++i) { a = (a + p[i]) % 65521u; b = (b + a) % 65521u; }
        return (b << 16) | a;
    }
    // zlib stream made of stored blocks
    static Bytes zlib_stored(const Bytes& in) {
        Bytes out = {0x78, 0x01};
        size_t pos = 0;
        do {
            const size_t len = std::min<size_t>(65535, in.size() - pos);
            const bool last = pos + len == in.size();
            out.push_back(last ? 1 : 0);
            out.push_back((uint8_t)(len & 255)); out.push_back((uint8_t)(len >> 8));
            out.push_back((uint8_t)(~len & 255)); out.push_back((uint8_t)((~len >> 8) & 255));
            out.insert(out.end(), in.begin() + (long)pos, in.begin() + (long)(pos + len));
            pos += len;
        } while (pos < in.size());
        const uint32_t ad = adler32(in.data(), in.size());
        for (int s = 24; s >= 0; s -= 8) out.push_back((uint8_t)(ad >> s));
        return out;
    }
    // chunked dataset: chunks of at most 8 rows (rank 2) / 16 elements (rank 1), one leaf node of a version-1 chunk B-tree
    Child dataset_chunked(const std::vector<size_t>& dims, const std::vector<float>& data) {
        const size_t rank = dims.size();
        std::vector<size_t> cd(rank);
        if (rank == 1) cd[0] = std::min<size_t>(dims[0], 16);
        else { cd[0] = std::min<size_t>(dims[0], 8); for (size_t k = 1; k < rank; ++k) cd[k] = dims[k]; }
        size_t chunk_elems = 1;
        for (size_t v : cd) chunk_elems *= v;
        const size_t n_chunks = (dims[0] + cd[0] - 1) / cd[0];
        if (n_chunks > 2 * 32) throw FormatError("too many chunks for a one-node chunk B-tree");
        const bool filtered = storage_ == 2;
        struct Stored { uint64_t addr; uint32_t size; size_t row0; };
        std::vector<Stored> stored;
        size_t inner = 1;
        for (size_t k = 1; k < rank; ++k) inner *= dims[k];
        for (size_t c = 0; c < n_chunks; ++c) {
            Bytes raw(chunk_elems * 4, 0);
            const size_t row0 = c * cd[0], rows = std::min(cd[0], dims[0] - row0);
            std::memcpy(raw.data(), data.data() + row0 * inner, rows * inner * 4);
            if (filtered) {
                Bytes sh(raw.size());
                const size_t cnt = raw.size() / 4;
                for (size_t e = 0; e < cnt; ++e)
                    for (unsigned b = 0; b < 4; ++b) sh[(size_t)b * cnt + e] = raw[e * 4 + b];
                raw = zlib_stored(sh);
            }
            const uint64_t at = alloc(raw.size());
            std::memcpy(buf_.data() + at, raw.data(), raw.size());
            stored.push_back({at, (uint32_t)raw.size(), row0});
        }
        // B-tree node: "TREE", type 1, level 0, entries, siblings undefined, then key/child pairs and the final key
        const size_t key = 8 + 8 * (rank + 1);
        const uint64_t node = alloc(24 + (2 * 32 + 1) * key + 2 * 32 * 8);
        std::memcpy(buf_.data() + node, "TREE", 4);
        put(node + 4, 1, 1); put(node + 5, 0, 1); put(node + 6, stored.size(), 2);
        put(node + 8, UNDEF, 8); put(node + 16, UNDEF, 8);
        uint64_t p = node + 24;
        for (const Stored& st : stored) {
            put(p, st.size, 4); put(p + 4, 0, 4);
            put(p + 8, st.row0, 8);
            for (size_t k = 1; k <= rank; ++k) put(p + 8 + 8 * k, 0, 8);
            put(p + key, st.addr, 8);
            p += key + 8;
        }
        put(p, 0, 4); put(p + 4, 0, 4); put(p + 8, n_chunks * cd[0], 8);    // final key: one past the last chunk
        for (size_t k = 1; k <= rank; ++k) put(p + 8 + 8 * k, 0, 8);
        Bytes space, type, fill, layout, pipeline;
        app(space, 1, 1); app(space, dims.size(), 1); app(space, 1, 1); app(space, 0, 5);
        for (size_t d : dims) app(space, d, 8);
        for (size_t d : dims) app(space, d, 8);
        app(type, 0x11, 1); app(type, 0x20, 1); app(type, 0x1f, 1); app(type, 0, 1); app(type, 4, 4);
        app(type, 0, 2); app(type, 32, 2); app(type, 23, 1); app(type, 8, 1); app(type, 0, 1); app(type, 23, 1); app(type, 127, 4);
        app(fill, 2, 1); app(fill, 2, 1); app(fill, 2, 1); app(fill, 1, 1); app(fill, 0, 4);
        app(layout, 3, 1); app(layout, 2, 1); app(layout, rank + 1, 1); app(layout, node, 8);
        for (size_t v : cd) app(layout, v, 4);
        app(layout, 4, 4);
        std::vector<Bytes> msgs = {message(0x0001, 0, space), message(0x0003, 1, type), message(0x0005, 1, fill)};
        if (filtered) {   // filter pipeline v1: shuffle (id 2, one client value: element size), deflate (id 1, one: level)
            app(pipeline, 1, 1); app(pipeline, 2, 1); app(pipeline, 0, 6);
            for (int f = 0; f < 2; ++f) {
                app(pipeline, f == 0 ? 2 : 1, 2); app(pipeline, 0, 2); app(pipeline, 1, 2); app(pipeline, 1, 2);
                app(pipeline, f == 0 ? 4 : 1, 4); app(pipeline, 0, 4);      // one client value + padding
            }
            msgs.push_back(message(0x000B, 0, pipeline));
        }
        msgs.push_back(message(0x0008, 0, layout));
        Child c;
        c.header = object_header(msgs);
        return c;
    }
    Child dataset(const std::vector<size_t>& dims, const std::vector<float>& data) {
        if (storage_ != 0 && !data.empty()) return dataset_chunked(dims, data);
        const uint64_t raw = alloc(data.size() * 4);
        if (!data.empty()) std::memcpy(buf_.data() + raw, data.data(), data.size() * 4);
        Bytes space, type, fill, layout;
        app(space, 1, 1); app(space, dims.size(), 1); app(space, 1, 1); app(space, 0, 5);
        for (size_t d : dims) app(space, d, 8);
        for (size_t d : dims) app(space, d, 8);                                          // maximum dimensions
        app(type, 0x11, 1); app(type, 0x20, 1); app(type, 0x1f, 1); app(type, 0, 1); app(type, 4, 4);
        app(type, 0, 2); app(type, 32, 2); app(type, 23, 1); app(type, 8, 1); app(type, 0, 1); app(type, 23, 1); app(type, 127, 4);
        app(fill, 2, 1); app(fill, 2, 1); app(fill, 2, 1); app(fill, 1, 1); app(fill, 0, 4);
        app(layout, 3, 1); app(layout, 1, 1); app(layout, raw, 8); app(layout, data.size() * 4, 8);
        Child c;
        c.header = object_header({message(0x0001, 0, space), message(0x0003, 1, type), message(0x0005, 1, fill),
                                  message(0x0008, 0, layout)});
        return c;
    }
    // children must be sorted by name (strcmp order), at most 2*LEAF_K * 2*INTERNAL_K of them
    Child group(const std::vector<std::pair<std::string, Child>>& children, const std::vector<Bytes>& attrs) {
        const size_t per_leaf = 2 * LEAF_K;
        const size_t n_leaves = children.empty() ? 0 : (children.size() + per_leaf - 1) / per_leaf;
        if (n_leaves > 2 * INTERNAL_K) throw FormatError("too many entries for a one-level group B-tree");
        // local heap: "" at offset 0, names 8-aligned, one trailing free block
        std::vector<uint64_t> name_off;
        Bytes seg(8, 0);
        for (const auto& c : children) {
            name_off.push_back(seg.size());
            seg.insert(seg.end(), c.first.begin(), c.first.end());
            seg.push_back(0);
            pad8(seg);
        }
        const uint64_t free_at = seg.size();
        const uint64_t seg_size = std::max<uint64_t>(88, free_at + 16);
        seg.resize(seg_size, 0);
        for (unsigned i = 0; i < 8; ++i) seg[free_at + i] = (uint8_t)(i == 0 ? 1 : 0);            // next free block: none (1)
        for (unsigned i = 0; i < 8; ++i) seg[free_at + 8 + i] = (uint8_t)((seg_size - free_at) >> (8 * i));
        Child g;
        g.is_group = true;
        g.btree = alloc(24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8);
        g.heap = alloc(32);
        const uint64_t seg_at = alloc(seg.size());
        std::memcpy(buf_.data() + seg_at, seg.data(), seg.size());
        std::memcpy(buf_.data() + g.heap, "HEAP", 4);
        put(g.heap + 8, seg_size, 8); put(g.heap + 16, free_at, 8); put(g.heap + 24, seg_at, 8);
        std::memcpy(buf_.data() + g.btree, "TREE", 4);
        put(g.btree + 4, 0, 1); put(g.btree + 5, 0, 1); put(g.btree + 6, n_leaves, 2);
        put(g.btree + 8, UNDEF, 8); put(g.btree + 16, UNDEF, 8);
        uint64_t kp = g.btree + 24;
        put(kp, 0, 8);                                                                        // key 0: the empty name
        for (size_t leaf = 0; leaf < n_leaves; ++leaf) {
            const size_t first = leaf * per_leaf, last = std::min(children.size(), first + per_leaf);
            const uint64_t snod = alloc(8 + per_leaf * 40);
            std::memcpy(buf_.data() + snod, "SNOD", 4);
            put(snod + 4, 1, 1); put(snod + 6, last - first, 2);
            for (size_t i = first; i < last; ++i) {
                const uint64_t e = snod + 8 + (i - first) * 40;
                const Child& c = children[i].second;
                put(e, name_off[i], 8); put(e + 8, c.header, 8);
                if (c.is_group) { put(e + 16, 1, 4); put(e + 24, c.btree, 8); put(e + 32, c.heap, 8); }
            }
            put(kp + 8, snod, 8);                                                             // child pointer
            put(kp + 16, name_off[last - 1], 8);                                              // key: largest name in the child
            kp += 16;
        }
        Bytes stab;
        app(stab, g.btree, 8); app(stab, g.heap, 8);
        std::vector<Bytes> msgs;
        msgs.push_back(message(0x0011, 0, stab));
        for (const Bytes& a : attrs) msgs.push_back(a);
        g.header = object_header(msgs);
        return g;
    }
};

inline void write_keras_weights(const std::string& path, const std::vector<KerasLayer>& layers, int storage = 0) {
    Writer::keras_weights(path, layers, storage);
}

}  // namespace h5
}  // namespace nsdf
#endif  // NSDF_H5_HPP_
