// The reference's own ProcessCirc::do_process, reachable from integration/gpu_circ_detect.cpp (which sees the class with that
// member renamed, see gpu_shim.h): CM_GPU_STAGE2=0 and single-end runs go here.
#include "process_circ.h"
void cm_ref_do_process(ProcessCirc *p) { p->do_process(); }
