// Auxiliar_functions.h -- forwarding header: the reference's header name, served by the B200-native host layer (see minipic_compat.h).
#ifndef MINIPIC_FWD_AUXILIAR_FUNCTIONS_H
#define MINIPIC_FWD_AUXILIAR_FUNCTIONS_H
#include "minipic_compat.h"
#endif
