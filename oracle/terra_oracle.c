#include "terra_oracle.h"
int terra_oracle_placeholder(void) { return 0; }
