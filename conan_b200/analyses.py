"""Stage 3 of the reference's script body (conan.py:1887-2090): Pearson, PCA, frame-frame clustering.

Filled in as the CUDA kernels for those contractions land; run_all dispatches on the keywords.
"""


def run_all(opts, ret, s2, trunc, trunc_inter, nres, nterm, asymm_data):
    pending = [k for k in ('pearson_time', 'pearson_obs_str', 'pca_make', 'pca_dir', 'k_traj_clusters', 'blocking',
                           'k_res_clusters', 'zoom') if opts.get(k)]
    for k in pending:
        print("The %s analysis is not available in this build yet." % k.upper())
