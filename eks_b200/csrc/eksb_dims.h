// (state dim N, observation dim P) pairs the library is instantiated for.
// The reference only ever uses N=3 with P in {4,6,8}; the rest
// cover north_star's "state dim 2-6".
#pragma once
#define EKSB_FOR_EACH_DIM(X) X(2, 4) X(3, 4) X(3, 6) X(3, 8) X(4, 6) X(6, 8)
