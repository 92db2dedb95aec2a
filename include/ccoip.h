/*
 * ccoip.h -- source compatibility with programs written against the reference, which take the protocol's
 * default ports from this header (reference: ccoip/public_include/ccoip.h:11-29).  Everything lives in pccl.h here.
 */
#ifndef CCOIP_H_
#define CCOIP_H_
#include "pccl.h"
#endif /* CCOIP_H_ */
