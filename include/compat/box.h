// Forwarding header: lets sources written against the reference's "box.h" build against the B200 core.
#pragma once
#include "rt_compat_cuda.h"
