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
// float32 / float64 datatypes, layout v3 contiguous and compact.  Anything else throws
// nsdf::h5::FormatError.
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

    void read_dataset(Node& d, const std::vector<Msg>& msgs) const {
        bool have_dims = false, have_layout = false;
        unsigned elem = 0;   // 4 = float32, 8 = float64
        uint64_t data_off = 0, data_size = 0;
        bool compact = false;
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
                } else {
                    throw FormatError("chunked dataset '" + d.name + "' is not supported");
                }
                have_layout = true;
            }
        }
        if (!have_dims || !have_layout) throw FormatError("dataset '" + d.name + "' lacks dataspace or layout");
        if (!elem) throw FormatError("dataset '" + d.name + "': only little-endian IEEE float32/float64 supported");
        uint64_t n = 1;
        for (size_t v : d.dims) {
            if (v != 0 && n > (uint64_t)buf_.size() / v) throw FormatError("dataset '" + d.name + "' larger than file");
            n *= v;
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
inline bool read_keras_layers(const File& f, std::vector<KerasLayer>& out, std::string& error) {
    out.clear();
    for (const auto& kv : f.root().children) {
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

}  // namespace h5
}  // namespace nsdf
#endif  // NSDF_H5_HPP_
