// Stand-in for <cpl/Utility.h> (see cpl/Common.h in this directory).
#pragma once
#include "Common.h"
