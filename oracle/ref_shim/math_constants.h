// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).  The two CUDA math_constants.h
// values the reference uses (published constants of the CUDA toolkit).
#pragma once
#define CUDART_PI_F   3.141592654f
#define CUDART_PIO2_F 1.570796327f
