// bv_loader.cpp -- see bv_loader.h.
#include "bv_loader.h"

#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <thread>

namespace wgbv {

static thread_local char t_error[512] = "";

int fail(int code, const char *fmt, ...) {
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(t_error, sizeof t_error, fmt, ap);
	va_end(ap);
	return code;
}
const char *last_error_message() { return t_error; }
void clear_error() { t_error[0] = 0; }

// ------------------------------------------------------------------------------------------------ properties
static std::string trim(const std::string &s) {
	size_t a = 0, b = s.size();
	while (a < b && (s[a] == ' ' || s[a] == '\t' || s[a] == '\r' || s[a] == '\f')) a++;
	while (b > a && (s[b - 1] == ' ' || s[b - 1] == '\t' || s[b - 1] == '\r' || s[b - 1] == '\f')) b--;
	return s.substr(a, b - a);
}

// java.util.Properties subset: "key=value", "key:value" or "key value" lines, '#' / '!' comments.
static void parse_kv(const char *text, std::map<std::string, std::string> &kv) {
	const char *p = text;
	while (*p) {
		const char *e = strchr(p, '\n');
		std::string line = e ? std::string(p, e) : std::string(p);
		p = e ? e + 1 : p + line.size();
		line = trim(line);
		if (line.empty() || line[0] == '#' || line[0] == '!') continue;
		size_t i = 0;
		while (i < line.size() && line[i] != '=' && line[i] != ':' && line[i] != ' ' && line[i] != '\t') i++;
		std::string key = line.substr(0, i);
		while (i < line.size() && (line[i] == ' ' || line[i] == '\t')) i++;
		if (i < line.size() && (line[i] == '=' || line[i] == ':')) i++;
		kv[key] = trim(line.substr(i < line.size() ? i : line.size()));
	}
}

static bool parse_i64(const std::map<std::string, std::string> &kv, const char *key, int64_t *out) {
	auto it = kv.find(key);
	if (it == kv.end() || it->second.empty()) return false;
	char *end = nullptr;
	const long long v = strtoll(it->second.c_str(), &end, 10);
	if (*end) return false;
	*out = v;
	return true;
}

// string2Flags (BVG:1316-1331): names are resolved against the constants BVGraph declares (BVG:475-524).
static int flag_value(const std::string &name, int *slot, int *coding) {
	static const struct { const char *name; int slot, coding; } table[] = {
		{"OUTDEGREES_GAMMA", 0, WGBV_GAMMA},   {"OUTDEGREES_DELTA", 0, WGBV_DELTA},   {"BLOCKS_GAMMA", 4, WGBV_GAMMA},
		{"BLOCKS_DELTA", 4, WGBV_DELTA},       {"RESIDUALS_GAMMA", 8, WGBV_GAMMA},    {"RESIDUALS_ZETA", 8, WGBV_ZETA},
		{"RESIDUALS_DELTA", 8, WGBV_DELTA},    {"RESIDUALS_NIBBLE", 8, WGBV_NIBBLE},  {"RESIDUALS_GOLOMB", 8, WGBV_GOLOMB},
		{"REFERENCES_GAMMA", 12, WGBV_GAMMA},  {"REFERENCES_DELTA", 12, WGBV_DELTA},  {"REFERENCES_UNARY", 12, WGBV_UNARY},
		{"BLOCK_COUNT_GAMMA", 16, WGBV_GAMMA}, {"BLOCK_COUNT_DELTA", 16, WGBV_DELTA}, {"BLOCK_COUNT_UNARY", 16, WGBV_UNARY},
		{"OFFSETS_GAMMA", 20, WGBV_GAMMA},     {"OFFSETS_DELTA", 20, WGBV_DELTA},
	};
	for (const auto &t : table)
		if (name == t.name) { *slot = t.slot; *coding = t.coding; return 1; }
	return 0;
}

int parse_properties_text(const char *text, wgbv_props *out) {
	if (!text || !out) return fail(WGBV_EARG, "null argument");
	std::map<std::string, std::string> kv;
	parse_kv(text, kv);
	// BVG:1491-1492: the class check accepts the standard (32-bit) BVGraph name too
	auto gc = kv.find("graphclass");
	std::string cls = gc == kv.end() ? "" : gc->second;
	if (cls.rfind("class ", 0) == 0) cls = cls.substr(6);
	if (cls != "it.unimi.dsi.big.webgraph.BVGraph" && cls != "it.unimi.dsi.webgraph.BVGraph")
		return fail(WGBV_EIO, "This class (GpuBVGraph) cannot load a graph stored using class \"%s\"", cls.c_str());
	int64_t v;
	if (!parse_i64(kv, "version", &v)) return fail(WGBV_EIO, "Missing format version information");
	if (v > 0) return fail(WGBV_EIO, "This graph uses format %lld, but this class can understand only graphs up to format 0", (long long)v);
	wgbv_props p;
	memset(&p, 0, sizeof p);
	p.outdegree_coding = WGBV_GAMMA;
	p.block_coding = WGBV_GAMMA;
	p.residual_coding = WGBV_ZETA;
	p.reference_coding = WGBV_UNARY;
	p.block_count_coding = WGBV_GAMMA;
	p.offset_coding = WGBV_GAMMA;
	p.zeta_k = 3;
	auto fl = kv.find("compressionflags");
	if (fl != kv.end() && !fl->second.empty()) {
		const std::string &s = fl->second;
		size_t a = 0;
		while (a <= s.size()) {
			size_t b = s.find('|', a);
			if (b == std::string::npos) b = s.size();
			const std::string name = trim(s.substr(a, b - a));
			int slot, coding;
			if (!flag_value(name, &slot, &coding)) return fail(WGBV_EIO, "Compression flag %s unknown.", name.c_str());
			switch (slot) {
			case 0: p.outdegree_coding = coding; break;
			case 4: p.block_coding = coding; break;
			case 8: p.residual_coding = coding; break;
			case 12: p.reference_coding = coding; break;
			case 16: p.block_count_coding = coding; break;
			default: p.offset_coding = coding;
			}
			a = b + 1;
		}
	}
	int64_t t;
	if (!parse_i64(kv, "nodes", &p.nodes) || p.nodes < 0) return fail(WGBV_EIO, "missing or malformed property: nodes");
	if (!parse_i64(kv, "arcs", &p.arcs)) return fail(WGBV_EIO, "missing or malformed property: arcs");
	if (!parse_i64(kv, "windowsize", &t) || t < 0 || t > 255) return fail(WGBV_EIO, "missing, malformed or unsupported property: windowsize");
	p.window_size = (int32_t)t;
	if (!parse_i64(kv, "maxrefcount", &t)) return fail(WGBV_EIO, "missing or malformed property: maxrefcount");
	p.max_ref_count = (int32_t)t;
	if (!parse_i64(kv, "minintervallength", &t) || t < 0) return fail(WGBV_EIO, "missing or malformed property: minintervallength");
	p.min_interval_length = (int32_t)t;
	if (parse_i64(kv, "zetak", &t)) {
		if (t < 1 || t > 32) return fail(WGBV_EIO, "unsupported zetak");
		p.zeta_k = (int32_t)t;
	}
	*out = p;
	return WGBV_OK;
}

int read_file(const std::string &path, std::vector<uint8_t> &out) {
	FILE *f = fopen(path.c_str(), "rb");
	if (!f) return fail(WGBV_EIO, "cannot open %s", path.c_str());
	fseek(f, 0, SEEK_END);
	const long long n = ftell(f);
	fseek(f, 0, SEEK_SET);
	out.resize((size_t)n);
	const size_t got = n ? fread(out.data(), 1, (size_t)n, f) : 0;
	fclose(f);
	if ((long long)got != n) return fail(WGBV_EIO, "short read on %s", path.c_str());
	return WGBV_OK;
}

// ------------------------------------------------------------------------------------------------ offsets
namespace {
struct HostBits {
	const uint8_t *p;
	uint64_t nbits, pos;
	bool overrun = false;
	uint64_t window() const {  // >= 57 valid bits, left aligned; zero beyond the end
		const uint64_t byte = pos >> 3;
		uint64_t w = 0;
		const uint64_t nbytes = (nbits + 7) >> 3;
		if (byte + 8 <= nbytes) {
			memcpy(&w, p + byte, 8);
			w = __builtin_bswap64(w);
		} else {
			for (int i = 0; i < 8; i++) w = (w << 8) | (byte + i < nbytes ? p[byte + i] : 0);
		}
		return w << (pos & 7);
	}
	uint64_t unary() {
		uint64_t z = 0;
		for (;;) {
			if (pos >= nbits) { overrun = true; return z; }
			const uint64_t w = window() | (1ull << 6);
			const int lz = __builtin_clzll(w);
			if (lz < 57) { pos += lz + 1; return z + lz; }
			pos += 57;
			z += 57;
		}
	}
	uint64_t bits(unsigned n) {
		uint64_t v = 0;
		while (n) {
			const unsigned take = n > 32 ? 32 : n;
			v = (v << take) | (window() >> (64 - take));
			pos += take;
			n -= take;
		}
		if (pos > nbits) overrun = true;
		return v;
	}
	uint64_t gamma() {
		// the whole code inside the window (at most 28 leading zeros): one load, one count, two shifts
		if (pos + 64 <= nbits) {
			uint64_t w;
			memcpy(&w, p + (pos >> 3), 8);
			w = __builtin_bswap64(w) << (pos & 7);
			if (w >> 35) {
				const int lz = __builtin_clzll(w);
				pos += 2 * (uint64_t)lz + 1;
				return ((w << lz) >> (63 - lz)) - 1;
			}
		}
		const uint64_t m = unary();
		if (m > 63) { overrun = true; return 0; }
		return ((1ull << m) | (m ? bits((unsigned)m) : 0)) - 1;
	}
	uint64_t delta() {
		const uint64_t m = gamma();
		if (m > 63) { overrun = true; return 0; }
		return ((1ull << m) | (m ? bits((unsigned)m) : 0)) - 1;
	}
};
}  // namespace

int decode_offsets(const wgbv_props &p, const uint8_t *bytes, uint64_t nbytes, int64_t count, std::vector<uint64_t> &out) {
	if (p.offset_coding != WGBV_GAMMA && p.offset_coding != WGBV_DELTA)
		return fail(WGBV_EUNSUPPORTED, "The required offset coding (%d) is not supported.", p.offset_coding);
	HostBits s{bytes, nbytes * 8, 0};
	out.resize((size_t)count);
	uint64_t acc = 0;
	const bool g = p.offset_coding == WGBV_GAMMA;
	for (int64_t i = 0; i < count; i++) {
		acc += g ? s.gamma() : s.delta();
		out[(size_t)i] = acc;
	}
	if (s.overrun) return fail(WGBV_EIO, "the offsets stream ends before %lld offsets were read", (long long)count);
	return WGBV_OK;
}

// ------------------------------------------------------------------------------------------------ writeOffsets
// BVGraph.writeOffsets (BVG:2595-2609) for a graph that comes without its .offsets file: the position of every record.
// Sequential by construction -- a record ends where its last code ends, and how many codes it has depends on the
// outdegree of the row it refers to, which sits at the start of an EARLIER record: no record can be parsed before the
// ones in front of it (the reference walks the file with one node iterator for the same reason).  So this is host
// code of the loader, like the parsing of .properties: one pass, lengths only, nothing is decoded into lists.
// Record layout: BVG:995-1097 (outdegree; reference; block count, blocks; interval count, left/length pairs; residuals).
namespace {
struct RecordSkipper : HostBits {
	uint64_t zeta(int k) {
		const uint64_t h = unary();
		if (h * (uint64_t)k + (uint64_t)k > 64) { overrun = true; return 0; }
		const uint64_t left = 1ull << (h * k);
		const uint64_t m = bits((unsigned)(h * k + k - 1));
		if (m < left) return m + left - 1;
		return ((m << 1) | bits(1)) - 1;
	}
	uint64_t nibble() {
		uint64_t x = 0;
		uint64_t stop;
		do {
			stop = bits(1);
			x = (x << 3) | bits(3);
		} while (!stop && !overrun);
		return x;
	}
	uint64_t golomb(int b) {
		if (b == 0) return 0;
		const uint64_t q = unary();
		int log2b = 0;
		while ((2ull << log2b) <= (uint64_t)b) log2b++;
		const uint64_t m = (1ull << (log2b + 1)) - (uint64_t)b;
		uint64_t x = log2b ? bits((unsigned)log2b) : 0;
		if (x >= m) x = ((x << 1) | bits(1)) - m;
		return q * (uint64_t)b + x;
	}
	uint64_t code(int coding, int k) {
		switch (coding) {
		case WGBV_GAMMA: return gamma();
		case WGBV_DELTA: return delta();
		case WGBV_UNARY: return unary();
		case WGBV_ZETA: return zeta(k);
		case WGBV_NIBBLE: return nibble();
		case WGBV_GOLOMB: return golomb(k);
		default: overrun = true; return 0;
		}
	}
};
}  // namespace

int build_offsets(const wgbv_props &p, const uint8_t *graph, uint64_t graph_bytes, std::vector<uint64_t> &out) {
	RecordSkipper s;
	s.p = graph;
	s.nbits = graph_bytes * 8;
	s.pos = 0;
	const int64_t n = p.nodes;
	const int w = p.window_size > 0 ? p.window_size : 0;
	std::vector<uint64_t> ring((size_t)w + 1, 0);  // outdegrees of the last w + 1 nodes
	out.resize((size_t)n + 1);
	for (int64_t x = 0; x < n; x++) {
		out[(size_t)x] = s.pos;
		const uint64_t d = s.code(p.outdegree_coding, 0);
		ring[(size_t)(x % (w + 1))] = d;
		if (d == 0) continue;
		if (d > 0x7fffffffull) return fail(WGBV_ESTATE, "node %lld: outdegree %llu", (long long)x, (unsigned long long)d);
		uint64_t ref = 0;
		if (w > 0) {
			ref = s.code(p.reference_coding, 0);
			if (ref > (uint64_t)w || (int64_t)ref > x) return fail(WGBV_ESTATE, "The required reference is incompatible with the window size (%d)", w);
		}
		uint64_t extra = d;
		if (ref) {
			const uint64_t dref = ring[(size_t)((x - (int64_t)ref) % (w + 1))];
			const uint64_t bc = s.code(p.block_count_coding, 0);
			uint64_t copied = 0, total = 0;
			for (uint64_t i = 0; i < bc && !s.overrun; i++) {
				const uint64_t len = s.code(p.block_coding, 0) + (i ? 1 : 0);
				total += len;
				if (!(i & 1)) copied += len;
			}
			if (!(bc & 1)) copied += dref > total ? dref - total : 0;  // the implicit last block copies the rest of the list
			if (copied > d) return fail(WGBV_ESTATE, "node %lld copies more successors than it has", (long long)x);
			extra = d - copied;
		}
		if (extra && p.min_interval_length != 0) {
			const uint64_t ic = s.gamma();
			for (uint64_t i = 0; i < ic && !s.overrun; i++) {
				s.gamma();  // left extreme
				const uint64_t len = s.gamma() + (uint64_t)p.min_interval_length;
				if (len > extra) return fail(WGBV_ESTATE, "node %lld: intervals longer than the successor list", (long long)x);
				extra -= len;
			}
		}
		for (uint64_t i = 0; i < extra && !s.overrun; i++) s.code(p.residual_coding, p.zeta_k);
		if (s.overrun || s.pos > s.nbits) return fail(WGBV_EIO, "the graph stream ends inside the record of node %lld", (long long)x);
	}
	out[(size_t)n] = s.pos;
	return WGBV_OK;
}

// the offsets as the .offsets stream codes them (gamma or delta of the differences, n + 1 codes, the first one 0)
void encode_offsets(const wgbv_props &p, const std::vector<uint64_t> &off, std::vector<uint8_t> &out) {
	out.clear();
	uint64_t acc = 0;
	int fill = 0;  // bits of acc not yet written (< 8 between calls)
	auto put32 = [&](uint64_t v, int len) {  // len <= 32
		if (!len) return;
		acc = (acc << len) | (v & ((1ull << len) - 1));
		fill += len;
		while (fill >= 8) {
			out.push_back((uint8_t)(acc >> (fill - 8)));
			fill -= 8;
		}
		acc &= (1ull << fill) - 1;
	};
	auto put = [&](uint64_t v, int len) {  // len <= 64
		if (len > 32) {
			put32(v >> 32, len - 32);
			len = 32;
		}
		put32(v, len);
	};
	auto put_gamma = [&](uint64_t x) {
		const uint64_t y = x + 1;
		const int m = 63 - __builtin_clzll(y);
		put(0, m);
		put(1, 1);
		put(y, m);  // the m bits below the leading one
	};
	uint64_t last = 0;
	for (size_t i = 0; i < off.size(); i++) {
		const uint64_t d = i ? off[i] - last : off[0];
		last = off[i];
		if (p.offset_coding == WGBV_DELTA) {
			const uint64_t y = d + 1;
			const int m = 63 - __builtin_clzll(y);
			put_gamma((uint64_t)m);
			put(y, m);
		} else put_gamma(d);
	}
	if (fill) out.push_back((uint8_t)(acc << (8 - fill)));
}

// ------------------------------------------------------------------------------------------------ Elias-Fano
// l = floor(log2(u / n)) lower bits (the choice of the reference's succinct lists, cf. EFGraph.java:140-142 and
// algo/EliasFanoCumulativeOutdegreeList.java:95-105); element i sets bit (v_i >> l) + i of the upper array.
// One hint per 64 elements; blocks whose ones are spread over more than 4096 bits keep explicit positions.
void EliasFano::build(const uint64_t *values, uint64_t n, uint64_t base) {
	count = n;
	lower.clear(); upper.clear(); hints.clear(); sparse.clear();
	if (n == 0) { l = 0; upper.assign(2, 0); hints.assign(1, 0); lower.assign(2, 0); sparse.assign(1, 0); return; }
	const uint64_t u = values[n - 1] - base + 1;
	l = 0;
	if (u / n >= 2) l = 63 - __builtin_clzll(u / n);
	const uint64_t upper_bits = n + ((values[n - 1] - base) >> l) + 1;
	upper.assign((size_t)((upper_bits + 63) / 64 + 2), 0);
	lower.assign((size_t)((n * (uint64_t)l + 63) / 64 + 2), 0);
	hints.assign((size_t)((n + 63) / 64 + 1), 0);
	const uint64_t lmask = l ? ((1ull << l) - 1) : 0;
	// Blocks of 64 elements are independent except for the words two neighbouring ranges share: the lower bits of 64
	// elements fill whole words (64 * l bits), the upper-bits words at the seams are merged with atomic ORs.  Sparse
	// blocks (rare: a giant record inside) are collected per range and numbered afterwards.
	const uint64_t nblocks = (n + 63) / 64;
	unsigned nt = std::thread::hardware_concurrency();
	if (nt == 0) nt = 1;
	if (nt > 16) nt = 16;
	if (nblocks < 4096) nt = 1;
	struct Sparse { uint64_t block; uint64_t pos[64]; };
	std::vector<std::vector<Sparse>> found(nt);
	auto work = [&](unsigned t) {
		const uint64_t b0 = nblocks * t / nt, b1 = nblocks * (t + 1) / nt;
		uint64_t cur_word = ~0ull, cur_bits = 0;
		auto flush = [&]() {
			if (cur_word != ~0ull) __atomic_fetch_or(&upper[(size_t)cur_word], cur_bits, __ATOMIC_RELAXED);
		};
		for (uint64_t k = b0; k < b1; k++) {
			const uint64_t i0 = k * 64, i1 = std::min(n, i0 + 64);
			uint64_t blockpos[64];
			for (uint64_t i = i0; i < i1; i++) {
				const uint64_t v = values[i] - base;
				const uint64_t pos = (v >> l) + i;
				if ((pos >> 6) != cur_word) { flush(); cur_word = pos >> 6; cur_bits = 0; }
				cur_bits |= 1ull << (pos & 63);
				if (l) {
					const uint64_t lb = i * (uint64_t)l;
					lower[(size_t)(lb >> 6)] |= (v & lmask) << (lb & 63);
					if ((lb & 63) + l > 64) lower[(size_t)(lb >> 6) + 1] |= (v & lmask) >> (64 - (lb & 63));
				}
				blockpos[i - i0] = pos;
			}
			const uint64_t first = blockpos[0], last = blockpos[i1 - i0 - 1];
			if (last - first > 4096) {
				Sparse sp;
				sp.block = k;
				for (uint64_t j = 0; j < 64; j++) sp.pos[j] = j < i1 - i0 ? blockpos[j] : last;
				found[t].push_back(sp);
			} else hints[(size_t)k] = first;
		}
		flush();
	};
	if (nt == 1) work(0);
	else {
		std::vector<std::thread> th;
		for (unsigned t = 0; t < nt; t++) th.emplace_back(work, t);
		for (auto &x : th) x.join();
	}
	for (auto &f : found)
		for (auto &sp : f) {
			hints[(size_t)sp.block] = (1ull << 63) | (uint64_t)(sparse.size() / 64);
			sparse.insert(sparse.end(), sp.pos, sp.pos + 64);
		}
	if (sparse.empty()) sparse.assign(1, 0);
}

static inline int select64_host(uint64_t x, int k) {
	for (; k > 0; k--) x &= x - 1;
	return __builtin_ctzll(x);
}

uint64_t EliasFano::get(uint64_t i) const {
	const uint64_t hint = hints[(size_t)(i >> 6)];
	const int r = (int)(i & 63);
	uint64_t p;
	if (hint >> 63) p = sparse[(size_t)(((hint & ~(1ull << 63)) << 6) + r)];
	else {
		uint64_t wi = hint >> 6;
		uint64_t word = upper[(size_t)wi] & (~0ull << (hint & 63));
		int need = r;
		for (;;) {
			const int c = __builtin_popcountll(word);
			if (need < c) break;
			need -= c;
			word = upper[(size_t)++wi];
		}
		p = (wi << 6) + select64_host(word, need);
	}
	const uint64_t high = p - i;
	uint64_t low = 0;
	if (l) {
		const uint64_t lb = i * (uint64_t)l;
		low = lower[(size_t)(lb >> 6)] >> (lb & 63);
		if ((lb & 63) + l > 64) low |= lower[(size_t)(lb >> 6) + 1] << (64 - (lb & 63));
		low &= (1ull << l) - 1;
	}
	return (high << l) | low;
}

// ------------------------------------------------------------------------------------------------ shards
void plan_shards(const std::vector<uint64_t> &offsets, int count, std::vector<int64_t> &bounds) {
	const int64_t n = (int64_t)offsets.size() - 1;
	bounds.assign((size_t)count + 1, n);
	bounds[0] = 0;
	const uint64_t total = offsets[(size_t)n];
	for (int k = 1; k < count; k++) {
		const uint64_t target = (uint64_t)((__uint128_t)total * (unsigned)k / (unsigned)count);
		const auto it = std::lower_bound(offsets.begin(), offsets.begin() + n, target);
		int64_t b = (int64_t)(it - offsets.begin());
		if (b < bounds[(size_t)k - 1]) b = bounds[(size_t)k - 1];
		bounds[(size_t)k] = b;
	}
}

}  // namespace wgbv
