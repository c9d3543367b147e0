// HDF5_output.h -- forwarding header: the reference's header name, served by the B200-native host layer (see minipic_compat.h).
#ifndef MINIPIC_FWD_HDF5_OUTPUT_H
#define MINIPIC_FWD_HDF5_OUTPUT_H
#include "minipic_compat.h"
#endif
