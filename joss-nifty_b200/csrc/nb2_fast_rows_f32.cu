#include "nb2_fast_impl.cuh"
namespace nb2 {
extern template int run_rows_part<float, 1>(nb2_plan*, const void*, void*, int, const nb2_pointop*, const nb2_pointop*, rt_stream);
extern template int run_rows_part<float, 2>(nb2_plan*, const void*, void*, int, const nb2_pointop*, const nb2_pointop*, rt_stream);
template int run_rows_part<float, 0>(nb2_plan*, const void*, void*, int, const nb2_pointop*, const nb2_pointop*, rt_stream);
template int run_rows<float>(nb2_plan*, const void*, void*, int, const nb2_pointop*, const nb2_pointop*, rt_stream);
template int rows_pairs_per_cta<float>(const nb2_plan*);
}
