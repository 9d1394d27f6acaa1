// decode.cu -- glue kernels of the 2B-4T-shaped decode step (filled in with the model harness).
#include "common.cuh"
