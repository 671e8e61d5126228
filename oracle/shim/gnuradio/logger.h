/* oracle/shim -- TEST INFRASTRUCTURE ONLY. No-op logger so fft_filter.cc compiles
 * without log4cpp (reference use: gr-filter/lib/fft_filter.cc:33,73). */
#ifndef GRB200_ORACLE_SHIM_LOGGER_H
#define GRB200_ORACLE_SHIM_LOGGER_H
#include <string>
namespace gr {
typedef void* logger_ptr;
static inline bool configure_default_loggers(logger_ptr& l, logger_ptr& d, const std::string&)
{
    l = nullptr;
    d = nullptr;
    return true;
}
} // namespace gr
#define GR_LOG_ALERT(log, msg) ((void)(log))
#define GR_LOG_DEBUG(log, msg) ((void)(log))
#define GR_LOG_INFO(log, msg) ((void)(log))
#define GR_LOG_WARN(log, msg) ((void)(log))
#define GR_LOG_ERROR(log, msg) ((void)(log))
#endif
