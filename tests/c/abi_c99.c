/* The drop-in boundary is a C ABI: this file must compile as strict C99 against include/kfslam_frontend.h (no C++-isms in the
 * header) and link against the library.  It calls only entry points that need no GPU: parameter defaults, the POD layouts the
 * C++ adapter relies on (28-byte cv::KeyPoint, 16-byte cv::DMatch), argument checking and the error string. */
#include <stdio.h>
#include <string.h>
#include "kfslam_frontend.h"

int main(void)
{
    kfs_orb_params p;
    kfs_replay_params rp;
    kfs_orb *orb = NULL;
    kfs_orb_default_params(&p);
    kfs_replay_default_params(&rp);
    if (sizeof(kfs_keypoint) != 28 || sizeof(kfs_dmatch) != 16) return 2;
    /* the reference's constructor arguments (VisualOdometry.cpp:10-16) except nfeatures, which is a run-time parameter */
    if (p.scale_factor != 1.2f || p.nlevels != 8 || p.edge_threshold != 31 || p.first_level != 0 || p.wta_k != 2 ||
        p.score_type != KFS_ORB_HARRIS_SCORE || p.patch_size != 31 || p.fast_threshold != 20) return 3;
    if (rp.ratio != 0.6f) return 4; /* VisualOdometry.cpp:46 */
    if (kfs_orb_create(NULL, 0, &orb) == KFS_OK) return 5; /* NULL parameters are an argument error with or without a GPU */
    if (strlen(kfs_last_error()) == 0) return 6;
    printf("abi_c99 ok: %d device(s)\n", kfs_device_count());
    return 0;
}
