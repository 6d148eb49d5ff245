#define PANN_INST_DT 1
#include "beam_search_inst.cuh"
