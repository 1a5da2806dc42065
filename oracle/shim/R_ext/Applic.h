/* shim header: see minir.h (test infrastructure only). */
#include "../minir.h"
