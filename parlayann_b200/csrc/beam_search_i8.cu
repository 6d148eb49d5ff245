#define PANN_INST_DT 2
#include "beam_search_inst.cuh"
