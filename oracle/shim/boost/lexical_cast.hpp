/* Stand-in for <boost/lexical_cast.hpp> (Boost is not in the image): the
 * reference's ParamsList only round-trips numbers through std::string
 * (Microchannel3D/.../params_list.h:31-33,118-121).  17 significant digits make
 * the double round trip lossless, as boost::lexical_cast does.  Test
 * infrastructure only. */
#ifndef LBM_ORACLE_SHIM_LEXICAL_CAST_HPP
#define LBM_ORACLE_SHIM_LEXICAL_CAST_HPP
#include <sstream>
#include <string>
#include <stdexcept>
namespace boost {
template <typename Target, typename Source>
Target lexical_cast(const Source &src) {
    std::stringstream ss;
    ss.precision(17);
    Target out;
    if (!(ss << src) || !(ss >> out)) throw std::runtime_error("bad lexical_cast");
    return out;
}
template <>
inline std::string lexical_cast<std::string, std::string>(const std::string &src) { return src; }
}  // namespace boost
#endif
