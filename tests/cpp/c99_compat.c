/* The public headers must be usable from plain C99 (reference: tests/c99_compat/enforce_c99_compat.c). */
#include <stdio.h>

#include "pccl.h"
#include "pccl_kernels.h"

int main(void) {
    pcclCommCreateParams_t params;
    pcclReduceDescriptor_t desc;
    pcclSharedState_t state;
    pcclKernelP2PArgs_t kargs;
    pcclBuildInfo_t info;
    (void) params;
    (void) desc;
    (void) state;
    (void) kargs;
    if (pcclInit() != pcclSuccess) return 1;
    if (pcclGetBuildInfo(&info) != pcclSuccess) return 1;
    if (pcclExtDataTypeSize(pcclBFloat16) != 2) return 1;
    if (sizeof(ccoip_socket_address_t) < 26) return 1;
    printf("c99 ok, cuda support: %d\n", (int) info.has_cuda_support);
    return 0;
}
