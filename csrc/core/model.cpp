#include "model.h"
#include "common.h"
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <unistd.h>

namespace tclb {

Model *Model::load(const std::string &path) {
    void *h = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!h) fail("cannot load model library %s: %s", path.c_str(), dlerror());
    typedef const tclb_model_t *(*get_t)(void);
    get_t g = (get_t)dlsym(h, "tclb_get_model");
    if (!g) fail("%s does not export tclb_get_model", path.c_str());
    const tclb_model_t *d = g();
    if (d->abi_version != TCLB_ABI_VERSION)
        fail("model library %s has ABI %d, runtime expects %d (rebuild it)", path.c_str(), d->abi_version, TCLB_ABI_VERSION);
    Model *m = new Model();
    m->d = d;
    m->handle = h;
    m->path = path;
    return m;
}
Model::~Model() {}

static std::string self_dir() {
    Dl_info info;
    if (dladdr((void *)&self_dir, &info) && info.dli_fname) {
        std::string p = info.dli_fname;
        size_t a = p.find_last_of('/');
        if (a != std::string::npos) return p.substr(0, a);
    }
    return ".";
}
std::string Model::find_library(const std::string &model, const std::string &precision) {
    std::string fn = "libtclb_" + model + "_" + precision + ".so";
    std::vector<std::string> dirs;
    if (const char *e = getenv("TCLB_MODEL_PATH")) for (auto &d : split(e, ':')) dirs.push_back(d);
    dirs.push_back(self_dir());
    dirs.push_back(self_dir() + "/../_lib");
    dirs.push_back("tclb_b200/_lib");
    for (auto &d : dirs) {
        std::string p = d + "/" + fn;
        if (access(p.c_str(), R_OK) == 0) return p;
    }
    fail("model library %s not found (searched TCLB_MODEL_PATH, %s); build it with `python -m tclb_b200.build %s`",
         fn.c_str(), self_dir().c_str(), model.c_str());
}
#define FINDER(fn, tab, cnt)                                            \
    int Model::fn(const std::string &n) const {                         \
        for (int i = 0; i < d->cnt; i++) if (n == d->tab[i].name) return i; \
        return -1;                                                      \
    }
FINDER(find_setting, settings, n_settings)
FINDER(find_zonesetting, zonesettings, n_zonesettings)
FINDER(find_global, globals, n_globals)
FINDER(find_quantity, quantities, n_quantities)
FINDER(find_field, fields, n_fields)
FINDER(find_density, densities, n_densities)
FINDER(find_action, actions, n_actions)
FINDER(find_stage, stages, n_stages)
const tclb_nodetype_t *Model::find_nodetype(const std::string &n) const {
    for (int i = 0; i < d->n_nodetypes; i++) if (n == d->nodetypes[i].name) return &d->nodetypes[i];
    return nullptr;
}
const tclb_nodegroup_t *Model::find_nodegroup(const std::string &n) const {
    for (int i = 0; i < d->n_nodegroups; i++) if (n == d->nodegroups[i].name) return &d->nodegroups[i];
    return nullptr;
}
bool Model::has_option(const std::string &n) const {
    for (int i = 0; i < d->n_options; i++) if (n == d->options[i].name) return d->options[i].value != 0;
    return false;
}
}  // namespace tclb
