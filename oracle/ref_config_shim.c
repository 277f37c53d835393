/* TEST INFRASTRUCTURE ONLY.
 * The reference's worker reads its configuration with libconfig 1.5 (src/service/service.cpp:1407-1437:
 * cfg.readFile, cfg.lookup("key") as strings, cfg.lookup("rocksdb") as an array).  This shim is the same look-ups over
 * the reference's vendored C library (submodules/libconfig/lib), so that the product's own parser of that file format
 * (include/rs_config.hpp) is checked against the real thing. */
#include <string.h>

#include "libconfig.h"

/* value of a string setting, "" when absent; returns 1 when found */
int refcfg_lookup_string(const char* path, const char* key, char* out, size_t cap) {
  config_t cfg;
  config_init(&cfg);
  int ok = 0;
  if (cap) out[0] = 0;
  if (config_read_file(&cfg, path)) {
    const char* v = 0;
    if (config_lookup_string(&cfg, key, &v) && v && strlen(v) + 1 <= cap) {
      strcpy(out, v);
      ok = 1;
    }
  } else {
    ok = -1; /* parse error */
  }
  config_destroy(&cfg);
  return ok;
}

/* elements of an array / list of strings joined by '\n'; returns their number, -1 on a parse error, -2 when absent */
int refcfg_lookup_array(const char* path, const char* key, char* out, size_t cap) {
  config_t cfg;
  config_init(&cfg);
  int n = -2;
  if (cap) out[0] = 0;
  if (!config_read_file(&cfg, path)) {
    n = -1;
  } else {
    config_setting_t* s = config_lookup(&cfg, key);
    if (s) {
      n = config_setting_length(s);
      size_t pos = 0;
      for (int i = 0; i < n; ++i) {
        const char* v = config_setting_get_string_elem(s, i);
        if (!v || pos + strlen(v) + 2 > cap) {
          n = -3;
          break;
        }
        strcpy(out + pos, v);
        pos += strlen(v);
        out[pos++] = '\n';
        out[pos] = 0;
      }
    }
  }
  config_destroy(&cfg);
  return n;
}
