#include "match.h"
