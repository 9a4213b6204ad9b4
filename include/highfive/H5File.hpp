// HighFive-subset shim: the part of HighFive 2.1 (a header-only wrapper over libhdf5, an un-vendored
// system library) that the reference calls in src/neuralNetwork.cpp:86-129 and src/simpleInfer.cpp:14-57,
// implemented over the library's own HDF5 reader (include/nsdf_h5.hpp).  Nothing here derives from
// HighFive or libhdf5 sources; only the names and call shapes are kept so the reference's files compile.
//
//   HighFive::File(path, HighFive::File::ReadOnly)          throws HighFive::Exception if unreadable
//   File / Group :: listObjectNames()                        ascending by name (H5_INDEX_NAME, H5_ITER_INC)
//                :: getObjectType(name) -> ObjectType::{Group, Dataset}
//                :: getGroup(name), getDataSet(name), getNumberObjects(), exist(name)
//   DataSet      :: getDimensions(), read(std::vector<float>&), read(std::vector<std::vector<float>>&)
#pragma once
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include "../nsdf_h5.hpp"

namespace HighFive {

class Exception : public std::runtime_error {
public:
    explicit Exception(const std::string& m) : std::runtime_error(m) {}
};
class FileException : public Exception { using Exception::Exception; };
class GroupException : public Exception { using Exception::Exception; };
class DataSetException : public Exception { using Exception::Exception; };

enum class ObjectType { File, Group, UserDataType, DataSpace, Dataset, Attribute, Other };

class DataSet;
class Group;

namespace details {
// Operations shared by File and Group (HighFive's NodeTraits).
class NodeView {
protected:
    std::shared_ptr<nsdf::h5::Node> node_;
    std::shared_ptr<nsdf::h5::Node> child(const std::string& name, const char* what) const {
        auto it = node_->children.find(name);
        if (it == node_->children.end()) throw GroupException(std::string("Unable to open ") + what + " \"" + name + "\"");
        return it->second;
    }

public:
    std::vector<std::string> listObjectNames() const {
        std::vector<std::string> names;
        for (const auto& kv : node_->children) names.push_back(kv.first);   // std::map: ascending byte order
        return names;
    }
    // Not HighFive: the names in the order of the `layer_names` attribute when it names exactly the objects present
    // (nsdf::h5::keras_layer_order), else ascending -- what include/neuralNetwork.inl walks so that stacks of more than
    // ten layers load in network order.
    std::vector<std::string> kerasLayerOrder() const { return nsdf::h5::keras_layer_order(*node_); }
    size_t getNumberObjects() const { return node_->children.size(); }
    bool exist(const std::string& name) const { return node_->children.count(name) != 0; }
    ObjectType getObjectType(const std::string& name) const {
        return child(name, "object")->kind == nsdf::h5::NodeKind::Group ? ObjectType::Group : ObjectType::Dataset;
    }
    inline Group getGroup(const std::string& name) const;
    inline DataSet getDataSet(const std::string& name) const;
};
}  // namespace details

class Group : public details::NodeView {
public:
    Group() {}
    explicit Group(std::shared_ptr<nsdf::h5::Node> n) { node_ = std::move(n); }
};

class File : public details::NodeView {
    std::shared_ptr<nsdf::h5::File> file_;
    std::string name_;

public:
    enum : unsigned { ReadOnly = 0x00u, ReadWrite = 0x01u, Truncate = 0x02u, Excl = 0x04u, Debug = 0x08u, Create = 0x10u,
                      Overwrite = Truncate, OpenOrCreate = ReadWrite | Create };
    explicit File(const std::string& filename, unsigned openFlags = ReadOnly) : name_(filename) {
        if (openFlags != ReadOnly) throw FileException("nsdf HighFive shim: only ReadOnly access is implemented");
        try {
            file_ = std::make_shared<nsdf::h5::File>(filename);
        } catch (const nsdf::h5::FormatError& e) {
            throw FileException(std::string("Unable to open file ") + filename + " (" + e.what() + ")");
        }
        node_ = file_->root_ptr();
    }
    const std::string& getName() const { return name_; }
};

}  // namespace HighFive

#include "H5DataSet.hpp"
