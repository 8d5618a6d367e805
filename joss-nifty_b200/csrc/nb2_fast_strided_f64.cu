#include "nb2_fast_impl.cuh"
namespace nb2 {
template int run_strided<double>(const nb2_plan*, int, const void*, void*, const nb2_pointop*, rt_stream,
                                 const RowRemap*);
}
