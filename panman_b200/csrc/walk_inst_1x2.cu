#include "walk_inst.cuh"
PMFS_WALK_INSTANCE(1, 2)
