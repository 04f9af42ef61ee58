// Source-compatibility forwarder: the reference header of this name (include/vq_bcn_compressor.h) is provided by
// include/yuptexture.h in this build.
#pragma once
#include "yuptexture.h"
