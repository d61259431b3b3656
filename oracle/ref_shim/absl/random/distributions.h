#include "random.h"
