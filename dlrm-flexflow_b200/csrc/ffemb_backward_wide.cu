// ffemb_backward_wide.cu — the 64-bit-key instantiations of the sorted backward (tables holding
// 2^32 - 1 rows or more on one GPU), as their own translation unit so that they compile in parallel
// with the 32-bit-key ones.  Everything is in ffemb_backward.cu.
#define FFEMB_BACKWARD_WIDE_TU
#include "ffemb_backward.cu"
