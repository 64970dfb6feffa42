"""B200-native GetPreferredAllocation path for the k8s GPU-sharing device plugin.

The directory name carries the upstream repository name (with hyphens), so it is loaded
under the importable alias ``k8s_gpu_sharing_plugin_b200`` by ``__graft_entry__.load_package()``.
"""
