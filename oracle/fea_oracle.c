/* TEST INFRASTRUCTURE ONLY -- the oracle (see fea_oracle_impl.h for the citations).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * liboracle.so; the product path (fea_visualizer_b200/csrc -> libfeab.so) never does.
 * Parity: PINNED -- the fp32 instantiation is checked against the compiled reference sources
 * (oracle/_ref/libfea_ref.so, tests/test_oracle_pin.py) and against committed golden vectors that
 * the reference produced (tests/golden/). */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define REAL float
#define SUF 32
#include "fea_oracle_impl.h"
#undef REAL
#undef SUF

#define REAL double
#define SUF 64
#include "fea_oracle_impl.h"
#undef REAL
#undef SUF
