#include "nb2_fast_cluster_impl.cuh"
namespace nb2 {
template int run_strided_cluster<double>(const nb2_plan*, int, int, long long, long long, const void*, void*,
                                      const nb2_pointop*, rt_stream, bool*);
}
