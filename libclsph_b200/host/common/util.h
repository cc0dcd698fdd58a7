// Z-curve helpers of libclsph/common/util.h (host half): 10 bits per axis, x in bit 0.
#ifndef _UTIL_H_
#define _UTIL_H_

#include "util/cl_types.h"

namespace sphb_host {
inline cl_uint spread_bits_10(cl_uint v) {
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}
inline cl_uint gather_bits_10(cl_uint v) {
  v &= 0x09249249u;
  v = (v | (v >> 2)) & 0x030C30C3u;
  v = (v | (v >> 4)) & 0x0300F00Fu;
  v = (v | (v >> 8)) & 0x030000FFu;
  v = (v | (v >> 16)) & 0x000003FFu;
  return v;
}
}  // namespace sphb_host

// libclsph/common/util.h:41-62
inline cl_uint get_grid_index_z_curve(cl_uint in_x, cl_uint in_y, cl_uint in_z) {
  return sphb_host::spread_bits_10(in_x) | (sphb_host::spread_bits_10(in_y) << 1) |
         (sphb_host::spread_bits_10(in_z) << 2);
}

// libclsph/common/util.h:4-38
inline cl_uint uninterleave(cl_uint value) { return sphb_host::gather_bits_10(value); }
inline cl_uint3 get_cell_coords_z_curve(cl_uint index) {
  cl_uint3 c;
  c.s[0] = sphb_host::gather_bits_10(index);
  c.s[1] = sphb_host::gather_bits_10(index >> 1);
  c.s[2] = sphb_host::gather_bits_10(index >> 2);
  c.s[3] = 0;
  return c;
}

#endif
