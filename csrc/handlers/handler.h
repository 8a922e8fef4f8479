// Handlers: XML element -> behaviour.  Scheduling contract and lifecycle follow the reference
// (src/Handlers/vHandler.h:33-102 Now/Next/Prev, src/Handlers.h:12-85 init-on-construct /
// finish-on-last-release, src/Handlers/GenericAction.cpp:10-39 child execution rules); classes
// self-register in a factory keyed by the element name (src/HandlerFactory.h:10-16).
#pragma once
#include <cmath>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include "../host/solver.h"

namespace tclb {

enum { HANDLER_CALLBACK = 0x01, HANDLER_ACTION = 0x02, HANDLER_DESIGN = 0x04, HANDLER_CONTAINER = 0x08 };
enum { PAR_GET = 0, PAR_SET = 1, PAR_GRAD = 2, PAR_UPPER = 3, PAR_LOWER = 4, PAR_X = 5, PAR_Y = 6, PAR_Z = 7, PAR_T = 8 };

class Handler {
  public:
    virtual ~Handler() {}
    XmlNode *node = nullptr;
    Solver *solver = nullptr;
    int start_iter = 0;
    double every_iter = 0;
    virtual int init() = 0;
    virtual int do_it() { return 0; }
    virtual int finish() { return 0; }
    virtual int type() const { return 0; }
    virtual int number_of_parameters() { return 0; }
    virtual int parameters(int /*what*/, double * /*data*/) { return 0; }
    bool now(double it) const {
        if (every_iter == 0) return false;
        it -= start_iter;
        return std::floor(it / every_iter) > std::floor((it - 1) / every_iter);
    }
    int next(double it) const {
        if (every_iter == 0) return -1;
        it -= start_iter;
        int k = (int)std::floor(it / every_iter);
        return (int)(-std::floor(-(k + 1) * every_iter) - it);
    }
    int prev(double it) const {
        if (every_iter == 0) return -1;
        it -= start_iter;
        int k = (int)std::floor((it - 1) / every_iter);
        return (int)(it + std::floor(-k * every_iter));
    }
    // attribute looked up on this element or the nearest ancestor that has it
    std::string context_attr(const std::string &name, bool *found = nullptr) const;
    void read_iterations();          // Iterations= attribute (unit aware) -> start_iter/every_iter
};

typedef std::shared_ptr<Handler> HandlerPtr;

class HandlerFactory {
  public:
    typedef std::function<Handler *()> Maker;
    static std::map<std::string, Maker> &table();
    static bool add(const std::string &xmlname, Maker m);
    // create + init; nullptr when init fails; unknown element -> error
    static HandlerPtr produce(XmlNode *node, Solver *solver);
};
#define TCLB_REGISTER_HANDLER(xmlname, cls) \
    static bool reg_##cls = ::tclb::HandlerFactory::add(xmlname, []() -> ::tclb::Handler * { return new cls(); });

// shared bases
class Callback : public Handler {
  public:
    int init() override { read_iterations(); return 0; }
    int type() const override { return HANDLER_CALLBACK; }
};
class Action : public Handler {
  public:
    int init() override;
    int type() const override { return HANDLER_ACTION; }
};
class GenericAction : public Action {
  public:
    int stack = 0;
    int init() override { stack = 0; return Action::init(); }
    int execute_internal();          // instantiate children
    void unstack();
    int finish() override;
    // common main loop of Solve / RunAction
    int run_loop(std::function<void(int steps, int iter_type, bool last)> step_fn);
};
class GenericContainer : public GenericAction {
  public:
    int init() override { GenericAction::init(); return execute_internal(); }
    int finish() override { unstack(); return GenericAction::finish(); }
};
class Design : public Handler {
  public:
    int init() override { read_iterations(); return 0; }
    int type() const override { return HANDLER_DESIGN; }
};

}  // namespace tclb
