// tclb: command-line driver.
//   tclb [options] <model> <case.xml> [XPATH op value ...] [-s file.xml] [-x]
// One executable serves every model (the model is a shared library picked at run time), otherwise the
// flow mirrors the reference's per-model `CLB/<model>/main case.xml ...` (src/main.cpp:173-425).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "core/common.h"
#include "host/solver.h"

using namespace tclb;

static void usage() {
    fprintf(stderr,
            "Usage: tclb [options] <model> <case.xml> [XPATH (=|@attr =|inject|insert|delete|print) value ...] [-s out.xml] [-x]\n"
            "Options:\n"
            "  --precision float|double|float_half|float_halfshift|double_float|double_floatshift\n"
            "                             arithmetic[_storage] of the model library (default float); the storage suffix is the\n"
            "                             reference's --with-storage: half = 16-bit lattice, shift = store value minus the field's shift\n"
            "  --cpu | --gpu              backend (default gpu)\n"
            "  --ranks N                  number of subdomains (y/z decomposition, one GPU each)\n"
            "  --devices a,b,c            GPU ids for the subdomains (default: round-robin over visible GPUs)\n"
            "  --halo direct|copy         fused in-kernel halo (default) or memcpy exchange baseline\n"
            "  --xblock N                 pad nx to a multiple of N (default 32 on GPU, 1 on CPU)\n"
            "  --lib path.so              explicit model library\n"
            "  --verbose L                console level 0..9 (default 3)\n");
}

int main(int argc, char **argv) {
    SolverOptions opt;
    int verbose = L_OUTPUT;
    std::vector<std::string> pos;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto need = [&](const char *what) -> std::string {
            if (i + 1 >= argc) { fprintf(stderr, "missing value for %s\n", what); exit(2); }
            return argv[++i];
        };
        if (pos.size() >= 2) { pos.push_back(a); continue; }      // everything after the case file is an XPATH edit
        if (a == "--precision") opt.precision = need("--precision");
        else if (a == "--simplepart") opt.simplepart_config = need("--simplepart");
        else if (a == "--no-graph") setenv("TCLB_NO_GRAPH", "1", 1);
        else if (a == "--cpu") opt.backend = TCLB_BACKEND_CPU;
        else if (a == "--gpu") opt.backend = TCLB_BACKEND_CUDA;
        else if (a == "--ranks") opt.nranks = atoi(need("--ranks").c_str());
        else if (a == "--devices") { for (auto &d : split(need("--devices"), ',')) opt.devices.push_back(atoi(d.c_str())); }
        else if (a == "--halo") opt.direct_halo = need("--halo") != "copy";
        else if (a == "--xblock") opt.x_block = atoi(need("--xblock").c_str());
        else if (a == "--lib") opt.model_path = need("--lib");
        else if (a == "--verbose") verbose = atoi(need("--verbose").c_str());
        else if (a == "-h" || a == "--help") { usage(); return 0; }
        else pos.push_back(a);
    }
    if (pos.size() < 2) { usage(); return 2; }
    log_init(0, verbose);
    opt.model = pos[0];
    std::vector<std::string> xargs(pos.begin() + 2, pos.end());
    logf(L_NOTICE_IMP, "-------------------------------------------------------------------------");
    logf(L_NOTICE_IMP, "-  tclb-b200   model: %-25s precision: %-6s          -", opt.model.c_str(), opt.precision.c_str());
    logf(L_NOTICE_IMP, "-------------------------------------------------------------------------");
    try {
        Solver solver(opt);
        solver.load_case(pos[1], xargs);
        return solver.run() == 0 ? 0 : 1;
    } catch (const std::exception &e) {
        logf(L_FATAL, "%s", e.what());
        return 1;
    }
}
