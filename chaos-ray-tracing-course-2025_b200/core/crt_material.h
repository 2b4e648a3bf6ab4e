#pragma once
// API-compatibility shim: in the reference this is src/core/crt_material.h; here the scene types live together.
#include "crt_scene.h"
