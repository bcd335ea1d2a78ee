// Part of the drop-in binary (oracle/_ref/circminer_gpu): the member gpu_shim.h renames mapping()'s
// genome_seeder.pac2char_whole_contig() to (src/circminer.cpp:244: the contig's text as one character per base, about a
// second per Gbp and round).  The drop-in maps from the 3-bit packed contig on the device, so a round does not need the text;
// cm_shim_unpack_contig (gpu_map_reads.cpp) decides.  Stage 2 unpacks a contig itself when one of its pairs needs the
// reference's own code (gpu_circ_detect.cpp).  Own translation unit: the class is declared with the renamed member here.
#define pac2char_whole_contig cm_shim_pac2char_whole_contig
#include "common.h"
#include "match_read.h"

bool cm_shim_unpack_contig();

bool GenomeSeeder::cm_shim_pac2char_whole_contig(void) { return cm_shim_unpack_contig(); }
