#include "orc_common.h"
