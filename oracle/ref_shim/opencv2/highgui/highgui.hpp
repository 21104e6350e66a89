// ORACLE — TEST INFRASTRUCTURE ONLY.  Stand-in for the OpenCV header of the same name (OpenCV's C++ headers are not
// installed in this image); see ../cv_ref_shim.h (or ../../cv_ref_shim.h) for what it provides and why.
#pragma once
#include "cv_ref_shim.h"
