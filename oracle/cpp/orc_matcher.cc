// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_matcher.h).
#include "orc_matcher.h"
