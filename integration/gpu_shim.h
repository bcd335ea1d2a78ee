/* Force-included (g++ -include) into the reference's src/circminer.cpp when building the drop-in binary
 * oracle/_ref/circminer_gpu.  The only change to the reference driver is where its mapping threads start:
 * mapping() calls pthread_create(..., map_reads, filter_args[th]) (src/circminer.cpp:290-293); here that call
 * is routed to cm_shim_pthread_create, which runs the batched GPU map_reads (integration/gpu_map_reads.cpp)
 * on thread 0 and nothing on the others.  Stage 2: the one call process_circ.do_process() of circ_detect (src/circminer.cpp:349-352)
 * is renamed to a member that integration/gpu_circ_detect.cpp defines (do_process with the single-split pairs batched through
 * libcm_gpu.so; CM_GPU_STAGE2=0 calls the reference's do_process).  Index loading: mapping()'s loadHashTable(&tmpTime)
 * (src/circminer.cpp:232) is renamed to cm_shim_loadHashTable (gpu_map_reads.cpp), which calls the reference's own
 * loadCompressedRefGenome -- the same index-file record with the k-mer table skipped, since the index is built on the device
 * (CM_GPU_REF_TABLE=1 keeps the reference's load).  The macro also renames the declaration in mrsfast/HashTable.h, which this
 * translation unit includes inside extern "C".  mapping()'s genome_seeder.pac2char_whole_contig() (:244, the contig as characters,
 * which only the reference's own extension reads) is renamed to a member that gpu_shim_members.cpp defines and that skips the
 * unpacking under the same switch.  Everything else -- FASTQ parsing, PAM / remain-FASTQ writers, the round loop
 * and the report of stage 2 -- is the reference's own code, unmodified. */
#pragma once
#include <pthread.h>
int cm_shim_pthread_create(pthread_t *t, const pthread_attr_t *a, void *(*fn)(void *), void *arg);
#define pthread_create cm_shim_pthread_create
#define do_process cm_shim_do_process
#define loadHashTable cm_shim_loadHashTable
#define pac2char_whole_contig cm_shim_pac2char_whole_contig
