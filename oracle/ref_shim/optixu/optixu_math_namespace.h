#pragma once
#include "../sutil/vec_math.h"
