/*
 * "path?key=value&key=value#cachefile" sugar of dmlc URIs, same outcome as the reference's
 * io::URISpec (src/io/uri_spec.h:43-76): at most one '#', at most one '?', every argument needs
 * both key and value, and with num_parts != 1 the cache name gains ".split<N>.part<i>".
 */
#ifndef DMLC_CORE_B200_URI_SPEC_H_
#define DMLC_CORE_B200_URI_SPEC_H_

#include <dmlc/logging.h>

#include <map>
#include <sstream>
#include <string>

namespace dmlc {
namespace b200 {

struct UriSpec {
  std::string uri;                           /* the path part */
  std::map<std::string, std::string> args;   /* ?k=v&k=v */
  std::string cache_file;                    /* #name (decorated), or empty */

  UriSpec(const std::string &raw, unsigned part_index, unsigned num_parts) {
    std::string rest = raw;
    const size_t hash = rest.find('#');
    if (hash != std::string::npos) {
      std::string cache = rest.substr(hash + 1);
      CHECK(cache.find('#') == std::string::npos)
          << "only one `#` is allowed in file path for cachefile specification";
      rest.resize(hash);
      if (num_parts != 1) {
        std::ostringstream os;
        os << cache << ".split" << num_parts << ".part" << part_index;
        cache = os.str();
      }
      cache_file = cache;
    }
    const size_t q = rest.find('?');
    if (q != std::string::npos) {
      const std::string query = rest.substr(q + 1);
      CHECK(query.find('?') == std::string::npos) << "only one `?` is allowed in a data URI";
      rest.resize(q);
      size_t pos = 0;
      int n = 0;
      while (pos <= query.size()) {
        size_t amp = query.find('&', pos);
        if (amp == std::string::npos) amp = query.size();
        const std::string kv = query.substr(pos, amp - pos);
        ++n;
        const size_t eq = kv.find('=');
        CHECK(!kv.empty() && eq != std::string::npos && eq != 0)
            << "Invalid uri argument format for key in arg " << n;
        CHECK(eq + 1 < kv.size()) << "Invalid uri argument format for value in arg " << n;
        args.insert(std::make_pair(kv.substr(0, eq), kv.substr(eq + 1)));
        pos = amp + 1;
        if (amp == query.size()) break;
      }
    }
    uri = rest;
  }
};

}  // namespace b200
}  // namespace dmlc
#endif  // DMLC_CORE_B200_URI_SPEC_H_
