// Handle on a loaded model library (dlopen of libtclb_<model>_<prec>.so).
// Plays the role of the reference's generated Model_m metadata object (src/Lists.h:13-160).
#pragma once
#include <map>
#include <string>
#include <vector>
#include "tclb_model_abi.h"

namespace tclb {

class Model {
  public:
    static Model *load(const std::string &path);           // throws tclb::Error
    static std::string find_library(const std::string &model, const std::string &precision);
    ~Model();
    const tclb_model_t *d;                                  // raw tables
    std::string path;
    int real_size() const { return d->real_size; }
    int flag_size() const { return d->flag_size; }
    int storage_size() const { return d->storage_size; }              // element size of a lattice value in memory
    double field_shift(int f) const { return d->storage_shifted ? d->field_shift[f] : 0.0; }
    int find_setting(const std::string &n) const;
    int find_zonesetting(const std::string &n) const;
    int find_global(const std::string &n) const;
    int find_quantity(const std::string &n) const;
    int find_field(const std::string &n) const;
    int find_density(const std::string &n) const;
    int find_action(const std::string &n) const;
    int find_stage(const std::string &n) const;
    const tclb_nodetype_t *find_nodetype(const std::string &n) const;
    const tclb_nodegroup_t *find_nodegroup(const std::string &n) const;
    bool has_option(const std::string &n) const;
    // bytes moved per node update by the reference's accounting (src/main.cpp:127)
    double bytes_per_update() const { return 2.0 * d->n_fields * d->storage_size + d->flag_size; }
  private:
    Model() : d(nullptr), handle(nullptr) {}
    void *handle;
};

}  // namespace tclb
