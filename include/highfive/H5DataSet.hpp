// HighFive-subset shim: DataSet (see H5File.hpp).
#pragma once
#include "H5File.hpp"

namespace HighFive {

class DataSet {
    std::shared_ptr<nsdf::h5::Node> node_;

public:
    DataSet() {}
    explicit DataSet(std::shared_ptr<nsdf::h5::Node> n) : node_(std::move(n)) {}
    std::vector<size_t> getDimensions() const { return node_->dims; }
    size_t getElementCount() const { return node_->data.size(); }

    void read(std::vector<float>& out) const {
        if (node_->dims.size() > 1) {   // HighFive squeezes only unit dimensions
            size_t big = 0;
            for (size_t d : node_->dims) big += d > 1;
            if (big > 1) throw DataSetException("Impossible to read DataSet of dimensions " + std::to_string(node_->dims.size()) + " into arrays of dimensions 1");
        }
        out = node_->data;
    }
    void read(std::vector<std::vector<float>>& out) const {
        if (node_->dims.size() != 2) throw DataSetException("Impossible to read DataSet of dimensions " + std::to_string(node_->dims.size()) + " into arrays of dimensions 2");
        const size_t r = node_->dims[0], c = node_->dims[1];
        out.assign(r, std::vector<float>(c));
        for (size_t i = 0; i < r; ++i)
            for (size_t j = 0; j < c; ++j) out[i][j] = node_->data[i * c + j];
    }
    void read(float& out) const {
        if (node_->data.size() != 1) throw DataSetException("Impossible to read a non-scalar DataSet into a scalar");
        out = node_->data[0];
    }
};

namespace details {
inline Group NodeView::getGroup(const std::string& name) const {
    auto c = child(name, "the group");
    if (c->kind != nsdf::h5::NodeKind::Group) throw GroupException("Unable to open the group \"" + name + "\"");
    return Group(c);
}
inline DataSet NodeView::getDataSet(const std::string& name) const {
    auto c = child(name, "the dataset");
    if (c->kind != nsdf::h5::NodeKind::Dataset) throw DataSetException("Unable to open the dataset \"" + name + "\"");
    return DataSet(c);
}
}  // namespace details

}  // namespace HighFive
