// oracle/ TBB stand-in (test infrastructure) - see tbbshim_core.h
#include "tbbshim_core.h"
