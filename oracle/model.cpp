/* model.cpp -- CPU restatement of a 2B-4T-shaped decode step built from the oracle
 * pieces (filled in with the decode harness). TEST INFRASTRUCTURE ONLY. */
#include "oracle.h"
