/* Force-included (g++ -include) into the reference's src/circminer.cpp when building the drop-in binary
 * oracle/_ref/circminer_gpu.  The only change to the reference driver is where its mapping threads start:
 * mapping() calls pthread_create(..., map_reads, filter_args[th]) (src/circminer.cpp:290-293); here that call
 * is routed to cm_shim_pthread_create, which runs the batched GPU map_reads (integration/gpu_map_reads.cpp)
 * on thread 0 and nothing on the others.  Stage 2: the one call process_circ.do_process() of circ_detect (src/circminer.cpp:349-352)
 * is renamed to a member that integration/gpu_circ_detect.cpp defines (do_process with the single-split pairs batched through
 * libcm_gpu.so; CM_GPU_STAGE2=0 calls the reference's do_process).  Everything else -- index loading, FASTQ parsing, PAM /
 * remain-FASTQ writers, the round loop, the double-split path and the report of stage 2 -- is the reference's own code, unmodified. */
#pragma once
#include <pthread.h>
int cm_shim_pthread_create(pthread_t *t, const pthread_attr_t *a, void *(*fn)(void *), void *arg);
#define pthread_create cm_shim_pthread_create
#define do_process cm_shim_do_process
