// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).
// ref_glue.cpp — the one stand-in file that includes the reference's own device headers: it
// defines the launch-parameter block every reference program declares
// (`extern "C" __constant__ SysParamter Parameter;`, e.g. raygen/raygen.cu:14) and tells the
// runtime where the path seed lives inside the reference's PerRayData (lib/raydata.cuh:60-80).
#include "shaders/sysparameter.h"

extern "C" {
SysParamter Parameter;
}

namespace refshim {
void* params_storage() { return &Parameter; }
unsigned int* prd_seed(unsigned int p0, unsigned int p1) { return &mergePointer(p0, p1)->seed; }
}
