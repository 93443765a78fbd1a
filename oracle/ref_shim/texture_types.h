// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).
#pragma once
typedef unsigned long long cudaTextureObject_t;
