#include "shim.h"
