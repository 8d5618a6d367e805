#include "nb2_fast_impl.cuh"
namespace nb2 {
template int run_rows_part<double, 1>(nb2_plan*, const void*, void*, int, const nb2_pointop*, const nb2_pointop*, rt_stream);
}
