/* oracle/shim -- TEST INFRASTRUCTURE ONLY. boost::make_unique -> std::make_unique
 * (reference use: gr-filter/lib/fft_filter.cc:18,92). */
#ifndef GRB200_ORACLE_SHIM_BOOST_MAKE_UNIQUE
#define GRB200_ORACLE_SHIM_BOOST_MAKE_UNIQUE
#include <memory>
namespace boost {
using std::make_unique;
}
#endif
