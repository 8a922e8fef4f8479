// Parameter protocol of the design elements (reference src/Handlers/vHandler.h:13-50, vHandler.cpp:30-66): the
// optimisation elements see ONE parameter vector, the concatenation over every design handler on the solver's stack.
#pragma once
#include "handler.h"

namespace tclb {

int design_count(Solver *s);                                 // design handlers currently on the stack
int design_number_of_parameters(Solver *s);
int design_parameters(Solver *s, int what, double *tab);     // what = PAR_GET | PAR_SET | PAR_GRAD | PAR_UPPER | PAR_LOWER | PAR_X..PAR_T
// all parameter fields on the DesignSpace nodes: the design used when the case file declares none
HandlerPtr make_default_design(Solver *s);

}  // namespace tclb
