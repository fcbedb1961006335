// oracle/shim/hdf5.h  --  TEST INFRASTRUCTURE.
//
// In-memory stand-in for the handful of HDF5 C calls the reference's measurement code makes
// (include/h5utils.h, include/measurementh5.h:276-361): H5Fcreate/H5Fclose, H5Gcreate2/H5Gclose,
// H5Screate_simple/H5Sclose, H5Dcreate2/H5Dwrite/H5Dclose.  HDF5 is absent from this image; with this
// header the reference's MeasurementManager compiles UNMODIFIED and every dataset it would have
// written (full path, dims, row-major payload) is recorded in h5shim::store(), which
// oracle/ref_measure.cpp exposes to the parity tests.  Only double datasets exist in the reference.
#pragma once
#include <sys/stat.h>
#include <sys/types.h>

#include <cstring>
#include <map>
#include <string>
#include <vector>

typedef long long hid_t;
typedef int herr_t;
typedef unsigned long long hsize_t;

#define H5P_DEFAULT 0
#define H5S_ALL 0
#define H5F_ACC_TRUNC 0u
#define H5T_NATIVE_DOUBLE 1

namespace h5shim {
struct Dataset {
    std::string path;
    std::vector<hsize_t> dims;
    std::vector<double> data;
};
struct State {
    std::vector<Dataset> sets;                      // in creation order
    std::map<hid_t, std::string> groups;            // file / group handle -> absolute path
    std::map<hid_t, std::vector<hsize_t>> spaces;   // dataspace handle -> dims
    std::map<hid_t, size_t> open_sets;              // dataset handle -> index into sets
    std::vector<std::string> files;                 // file names passed to H5Fcreate
    hid_t next = 100;
};
inline State& store() {
    static State s;
    return s;
}
inline std::string join(const std::string& parent, const char* name) {
    std::string n(name);
    if (!n.empty() && n[0] == '/') return n;  // absolute
    return parent + "/" + n;
}
}  // namespace h5shim

inline hid_t H5Fcreate(const char* name, unsigned, hid_t, hid_t) {
    auto& s = h5shim::store();
    s.files.push_back(name);
    const hid_t id = s.next++;
    s.groups[id] = "";
    return id;
}
inline herr_t H5Fclose(hid_t id) {
    h5shim::store().groups.erase(id);
    return 0;
}
inline hid_t H5Gcreate2(hid_t loc, const char* name, hid_t, hid_t, hid_t) {
    auto& s = h5shim::store();
    const hid_t id = s.next++;
    s.groups[id] = h5shim::join(s.groups[loc], name);
    return id;
}
inline herr_t H5Gclose(hid_t id) {
    h5shim::store().groups.erase(id);
    return 0;
}
inline hid_t H5Screate_simple(int rank, const hsize_t* dims, const hsize_t*) {
    auto& s = h5shim::store();
    const hid_t id = s.next++;
    s.spaces[id] = std::vector<hsize_t>(dims, dims + rank);
    return id;
}
inline herr_t H5Sclose(hid_t id) {
    h5shim::store().spaces.erase(id);
    return 0;
}
inline hid_t H5Dcreate2(hid_t loc, const char* name, hid_t, hid_t space, hid_t, hid_t, hid_t) {
    auto& s = h5shim::store();
    h5shim::Dataset d;
    d.path = h5shim::join(s.groups[loc], name);
    d.dims = s.spaces[space];
    size_t n = 1;
    for (hsize_t v : d.dims) n *= (size_t)v;
    d.data.assign(n, 0.0);
    s.sets.push_back(d);
    const hid_t id = s.next++;
    s.open_sets[id] = s.sets.size() - 1;
    return id;
}
inline herr_t H5Dwrite(hid_t dset, hid_t, hid_t, hid_t, hid_t, const void* buf) {
    auto& s = h5shim::store();
    auto it = s.open_sets.find(dset);
    if (it == s.open_sets.end()) return -1;
    h5shim::Dataset& d = s.sets[it->second];
    std::memcpy(d.data.data(), buf, d.data.size() * sizeof(double));
    return 0;
}
inline herr_t H5Dclose(hid_t id) {
    h5shim::store().open_sets.erase(id);
    return 0;
}
