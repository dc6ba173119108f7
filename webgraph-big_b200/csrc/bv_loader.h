// bv_loader.h -- host-side loading: .properties parsing, .offsets decoding, shard planning and the
// Elias-Fano construction (the host half of BVGraph.loadInternal, BVGraph.java:1479-1574 of the reference).
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/wgbv.h"

namespace wgbv {

// Sets the thread-local error message and returns `code`.
int fail(int code, const char *fmt, ...);
const char *last_error_message();
void clear_error();

int parse_properties_text(const char *text, wgbv_props *out);
int read_file(const std::string &path, std::vector<uint8_t> &out);

// `count` absolute bit offsets from the gamma/delta coded differences (OffsetsLongIterator, BVG:870-898).
int decode_offsets(const wgbv_props &p, const uint8_t *bytes, uint64_t nbytes, int64_t count, std::vector<uint64_t> &out);
// BVGraph.writeOffsets (BVG:2595-2609): the bit position of every record of a .graph stream, nodes + 1 values (host,
// sequential by construction); and the .offsets stream that codes them
int build_offsets(const wgbv_props &p, const uint8_t *graph, uint64_t graph_bytes, std::vector<uint64_t> &out);
void encode_offsets(const wgbv_props &p, const std::vector<uint64_t> &off, std::vector<uint8_t> &out);

// Elias-Fano representation of a monotone list, laid out for the device (see bv_device.cuh: ef_get).
struct EliasFano {
	int l = 0;
	uint64_t count = 0;
	std::vector<uint64_t> lower, upper, hints, sparse;
	void build(const uint64_t *values, uint64_t n, uint64_t base);
	uint64_t get(uint64_t i) const;  // host-side mirror of ef_get, used for accounting only
	uint64_t bytes() const { return 8 * (lower.size() + upper.size() + hints.size() + sparse.size()); }
};

// Contiguous shards balanced by compressed bits: bounds[k] is the first node of shard k, bounds[count] = n.
void plan_shards(const std::vector<uint64_t> &offsets, int count, std::vector<int64_t> &bounds);

}  // namespace wgbv
