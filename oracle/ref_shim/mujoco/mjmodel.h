// ORACLE BUILD SHIM (test infrastructure): see mujoco.h
#include "mujoco.h"
