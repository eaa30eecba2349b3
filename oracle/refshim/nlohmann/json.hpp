// Stand-in include path for nlohmann/json (single header shipped on the image
// by cudnn_frontend). Build-time only; used solely to compile the reference
// sources into oracle/_ref. The include directory is passed by the Makefile.
#pragma once
#include <cudnn_frontend/thirdparty/nlohmann/json.hpp>
