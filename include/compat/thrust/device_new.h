#pragma once
#include "../rt_compat_cuda.h"
