/* Stand-in for boost/utility/binary.hpp: the reference uses BOOST_BINARY only for the
 * top-bit masks in software_clz (R/Include/CLData/CLPortability.h:165,178,191).
 * TEST INFRASTRUCTURE for the oracle build only. */
#ifndef CLRT_ORACLE_SHIM_BOOST_BINARY
#define CLRT_ORACLE_SHIM_BOOST_BINARY
static inline unsigned long long clrt_boost_binary(const char* s) {
  unsigned long long v = 0;
  for (; *s; ++s) if (*s == '0' || *s == '1') v = (v << 1) | (unsigned long long)(*s - '0');
  return v;
}
#define BOOST_BINARY(x) (clrt_boost_binary(#x))
#endif
