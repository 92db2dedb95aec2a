/*
 * ccoip_inet.h -- source compatibility: the reference declares ccoip_socket_address_t and friends in a header of
 * this name (ccoip/public_include/ccoip_inet.h:13-35).  The same types, with the same layout, are part of pccl.h here.
 */
#ifndef CCOIP_INET_H_
#define CCOIP_INET_H_
#include "pccl.h"
#endif /* CCOIP_INET_H_ */
