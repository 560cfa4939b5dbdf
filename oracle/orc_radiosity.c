/* placeholder */
#include "osb_oracle.h"
