// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_primitives.h).  CPU restatement of the reference's ORBmatcher
// searches over plain arrays (src/ORBmatcher.cc).  Filled in by orc_matcher.cc.
#pragma once
#include <cstdint>
#include <vector>
