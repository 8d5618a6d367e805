#include "nb2_fast_impl.cuh"
namespace nb2 {
template int run_rows<double>(nb2_plan*, const void*, void*, int, const nb2_pointop*, const nb2_pointop*, rt_stream,
                          const RowGroup*);
template int rows_pairs_per_cta<double>(const nb2_plan*);
}
