#include "nb2_c2c_impl.cuh"
namespace nb2 {
template int run_rows_c2c<float>(const void*, void*, const void*, int, long long, int, int, double, rt_stream);
}
