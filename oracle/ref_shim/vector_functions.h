#pragma once
#include "vector_types.h"
