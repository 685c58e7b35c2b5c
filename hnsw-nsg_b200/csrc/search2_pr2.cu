// search kernel v2, pool form PR=2 (see search2_impl.cuh)
#define SEARCH2_PR 2
#include "search2_impl.cuh"
