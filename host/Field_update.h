// Field_update.h -- forwarding header: the reference's header name, served by the B200-native host layer (see minipic_compat.h).
#ifndef MINIPIC_FWD_FIELD_UPDATE_H
#define MINIPIC_FWD_FIELD_UPDATE_H
#include "minipic_compat.h"
#endif
