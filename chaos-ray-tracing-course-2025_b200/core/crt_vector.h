#pragma once
// API-compatibility shim: the reference splits this over src/core/crt_vector.h
#include "crt_math.h"
