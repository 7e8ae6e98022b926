// Process-wide count of kernels launched by this library (reported as gpu_launches by bench.py).
#pragma once
#include <stdint.h>
void bddl_count_launch(int64_t n);
