/* The boundary is a C ABI: this file is compiled as C99 (no C++), links libumx_cuda.so and calls the entry points that need no
 * GPU.  Prints the segment table of a 3-minute track shifted by the reference's default offset (umx.cpp:214-218). */
#include "umx_cuda.h"
#include <stdio.h>

int main(void)
{
    int64_t offsets[16];
    int lengths[16];
    const int n = umx_cuda_segment_plan(7938000, UMX_DEFAULT_SHIFT_OFFSET, UMX_SEGMENT_SAMPLES, offsets, lengths, 16);
    if (n < 0)
    {
        fprintf(stderr, "%s\n", umx_cuda_last_error());
        return 1;
    }
    printf("frames %d segments %d\n", umx_cuda_nb_frames(UMX_SEGMENT_SAMPLES), n);
    for (int i = 0; i < n; ++i)
        printf("%lld %d\n", (long long)offsets[i], lengths[i]);
    /* a null argument is an error return with a message, never a crash */
    if (umx_cuda_session_wait(NULL, 0) >= 0 || umx_cuda_last_error()[0] == 0)
        return 2;
    return 0;
}
