#define PANN_INST_DT 0
#include "beam_search_inst.cuh"
