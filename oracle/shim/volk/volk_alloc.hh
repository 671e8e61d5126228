/* oracle/shim/volk/volk_alloc.hh -- TEST INFRASTRUCTURE ONLY.
 * volk::vector stand-in (reference use: gr-filter/include/gnuradio/filter/fft_filter.h:18,60). */
#ifndef GRB200_ORACLE_SHIM_VOLK_ALLOC_HH
#define GRB200_ORACLE_SHIM_VOLK_ALLOC_HH
#include <vector>
namespace volk {
template <class T>
using vector = std::vector<T>;
}
#endif
