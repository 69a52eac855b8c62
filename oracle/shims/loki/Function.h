// TEST INFRASTRUCTURE ONLY: <loki/Function.h> pulls in Functor.h (see Singleton.h for why these shims exist)
#ifndef PSGPU_ORACLE_LOKI_FUNCTION_H
#define PSGPU_ORACLE_LOKI_FUNCTION_H
#include "Functor.h"
#endif
