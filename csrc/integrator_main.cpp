// bin/tclb-integrator: particle integrators as separate programs, the counterpart of <RemoteForceInterface spawn=/connect=>
// (csrc/host/rfi_bridge.h).  Roles of the reference's side programs: `simplepart` (src/simplepart.cpp) and `empty`
// (src/empty.cpp).
//   tclb-integrator simplepart <config.xml> [--log file.csv] [--omega] [--torque] (--fd N | --listen /path/to/socket)
//   tclb-integrator empty (--fd N | --listen /path/to/socket)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <memory>
#include <string>
#include "host/rfi_bridge.h"
#include "host/xml.h"

using namespace tclb;

int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s simplepart <config.xml> [--log f] [--omega] [--torque] (--fd N | --listen path)\n       %s empty (--fd N | --listen path)\n", argv[0], argv[0]); return 2; }
    std::string kind = argv[1], config, listen_path, logname = "SP_Log.csv";
    int fd = -1;
    bool rot = false, torque = false;
    for (int i = 2; i < argc; i++) {
        std::string a = argv[i];
        if (a == "--fd" && i + 1 < argc) fd = atoi(argv[++i]);
        else if (a == "--listen" && i + 1 < argc) listen_path = argv[++i];
        else if (a == "--log" && i + 1 < argc) logname = argv[++i];
        else if (a == "--omega") rot = true;
        else if (a == "--torque") torque = true;
        else if (config.empty()) config = a;
        else { fprintf(stderr, "unexpected argument %s\n", a.c_str()); return 2; }
    }
    std::unique_ptr<XmlDocument> doc;
    std::unique_ptr<ParticleIntegrator> integ;
    if (kind == "empty") integ.reset(new EmptyIntegrator());
    else if (kind == "simplepart") {
        if (config.empty()) { fprintf(stderr, "simplepart needs a configuration file\n"); return 2; }
        doc.reset(new XmlDocument());
        try { doc->load_file(config); } catch (const std::exception &e) { fprintf(stderr, "cannot read %s: %s\n", config.c_str(), e.what()); return 1; }
        const XmlNode *cfg = doc->root();
        if (cfg && cfg->name != "SimplePart") { const XmlNode *c = cfg->child("SimplePart"); if (c) cfg = c; }
        if (!cfg) { fprintf(stderr, "no <SimplePart> element in %s\n", config.c_str()); return 1; }
        integ.reset(new SimplePart(cfg, logname, rot, torque));
    } else { fprintf(stderr, "unknown integrator %s (simplepart, empty)\n", kind.c_str()); return 2; }
    if (fd < 0 && !listen_path.empty()) fd = rfi_listen_accept(listen_path);
    if (fd < 0) { fprintf(stderr, "no connection: give --fd N (spawned by the solver) or --listen path\n"); return 1; }
    long n = rfi_serve(fd, integ.get());
    if (n < 0) { fprintf(stderr, "handshake with the calculator failed\n"); return 1; }
    printf("tclb-integrator %s: served %ld exchanges\n", kind.c_str(), n);
    return 0;
}
