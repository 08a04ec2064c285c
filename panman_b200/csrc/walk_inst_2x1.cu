#include "walk_inst.cuh"
PMFS_WALK_INSTANCE(2, 1)
