// columns.cu — extra columns in HBM, predicate compiler, predicate kernel (see columns.h).
#include "columns.h"

#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace lb {

// ------------------------------------------------------------------------------------------------ kernels
// One thread per row, one warp per bitmap word.  The program is uniform across the grid (kernel parameter block),
// so every branch on an opcode is warp-uniform; a leaf is one coalesced 8-byte (or 4-byte) load per row plus one
// validity word per warp.  HBM-bound: 8 B/row per distinct column the predicate names (C5: 80 MB for 10 M rows),
// paid once per (predicate, epoch) — the result is cached on the device.
__device__ __forceinline__ bool pred_cmp_i64(int cmp, long long a, long long b) {
	return ((a < b ? 1 : (a > b ? 4 : 2)) & cmp) != 0;
}
// (a NaN is neither less nor greater: it lands on "equal", like the host mirror's evaluator)
__device__ __forceinline__ bool pred_cmp_f64(int cmp, double a, double b) {
	return ((a < b ? 1 : (a > b ? 4 : 2)) & cmp) != 0;
}

// PRED_U words (of 32 rows) per warp and iteration: the leaf loads of the PRED_U rows a thread owns are independent, so
// every warp keeps PRED_U x 256 B in flight per column.
static constexpr int PRED_U = 4;

// A leaf in two phases, so that the PRED_U rows of a thread have all their loads in flight together: (1) the validity
// word and the raw value are loaded unconditionally for rows inside the table (no load depends on another), (2) the
// comparison runs on the loaded bits.
__device__ __forceinline__ void pred_leaf_eval(const PredProgram &prog, const PredOp &o, bool in_range, uint32_t vword, int lane,
                                               long long raw, bool &t, bool &f) {
	t = false, f = false;
	if (!in_range) return;
	const bool valid = (vword >> lane) & 1u;
	if (o.op == PRED_IS_NULL) {
		t = !valid, f = valid;
	} else if (o.op == PRED_NOT_NULL) {
		t = valid, f = !valid;
	} else if (valid) {
		bool c = false, isnull = false;
		if (o.op == PRED_CMP_I64) {
			c = pred_cmp_i64(o.cmp, raw, o.ival);
		} else if (o.op == PRED_CMP_F64) {
			c = pred_cmp_f64(o.cmp, o.col_is_f64 ? __longlong_as_double(raw) : (double)raw, o.dval);
		} else if (o.op == PRED_IN_I64) {
			for (int j = 0; j < o.n_list; j++) c |= prog.aux_i64[o.aux_off + j] == raw;
			isnull = !c && o.has_null;
			if (o.negate) c = !c;
		} else if (o.op == PRED_IN_F64) {
			const double v = o.col_is_f64 ? __longlong_as_double(raw) : (double)raw;
			for (int j = 0; j < o.n_list; j++) c |= prog.aux_f64[o.aux_off + j] == v;
			isnull = !c && o.has_null;
			if (o.negate) c = !c;
		} else { // PRED_STR_LUT: raw = dictionary code
			const uint8_t v = prog.aux_u8[o.aux_off + raw];
			c = v == 1;
			isnull = v == 2;
		}
		t = c && !isnull;
		f = !c && !isnull;
	}
}

__global__ void __launch_bounds__(256) eval_predicate_kernel(const __grid_constant__ PredProgram prog, long long n,
                                                              uint32_t *__restrict__ out, long long words) {
	const long long warp0 = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
	const int lane = threadIdx.x & 31;
	for (long long wb = warp0 * PRED_U; wb < words; wb += nwarps * PRED_U) {
		uint32_t st[PRED_U], sf[PRED_U]; // per row: stacks of (true, false) bits, bit i = depth i
#pragma unroll
		for (int u = 0; u < PRED_U; u++) st[u] = 0u, sf[u] = 0u;
		int sp = 0;
		for (int i = 0; i < prog.n_ops; i++) {
			const PredOp &o = prog.ops[i];
			if (o.op == PRED_NOT) {
				const uint32_t m = 1u << (sp - 1);
#pragma unroll
				for (int u = 0; u < PRED_U; u++) {
					const uint32_t t = st[u] & m, f = sf[u] & m;
					st[u] = (st[u] & ~m) | f;
					sf[u] = (sf[u] & ~m) | t;
				}
			} else if (o.op == PRED_AND || o.op == PRED_OR) {
				const uint32_t ma = 1u << (sp - 2), mb = 1u << (sp - 1);
#pragma unroll
				for (int u = 0; u < PRED_U; u++) {
					const bool ta = st[u] & ma, fa = sf[u] & ma, tb = st[u] & mb, fb = sf[u] & mb;
					const bool t = o.op == PRED_AND ? (ta && tb) : (ta || tb);
					const bool f = o.op == PRED_AND ? (fa || fb) : (fa && fb);
					st[u] = (st[u] & ~(ma | mb)) | (t ? ma : 0u);
					sf[u] = (sf[u] & ~(ma | mb)) | (f ? ma : 0u);
				}
				sp--;
			} else {
				uint32_t vword[PRED_U];
				long long raw[PRED_U];
				const bool is_const = o.op == PRED_CONST, is_str = o.op == PRED_STR_LUT;
				const bool wants_value = !is_const && o.op != PRED_IS_NULL && o.op != PRED_NOT_NULL;
#pragma unroll
				for (int u = 0; u < PRED_U; u++) { // phase 1: every load of the PRED_U rows, none depends on another
					const long long w = wb + u, r = w * 32 + lane;
					const bool in = !is_const && w < words && r < n;
					vword[u] = in ? o.valid[w] : 0u;
					raw[u] = 0;
					if (in && wants_value)
						raw[u] = is_str ? (long long)static_cast<const int32_t *>(o.data)[r] : static_cast<const long long *>(o.data)[r];
				}
#pragma unroll
				for (int u = 0; u < PRED_U; u++) { // phase 2
					bool t, f;
					if (is_const) {
						t = o.ival == 1, f = o.ival == 0;
					} else {
						const long long w = wb + u, r = w * 32 + lane;
						pred_leaf_eval(prog, o, w < words && r < n, vword[u], lane, raw[u], t, f);
					}
					st[u] |= t ? (1u << sp) : 0u;
					sf[u] |= f ? (1u << sp) : 0u;
				}
				sp++;
			}
		}
#pragma unroll
		for (int u = 0; u < PRED_U; u++) {
			const long long w = wb + u, r = w * 32 + lane;
			const uint32_t word = __ballot_sync(0xffffffffu, w < words && r < n && (st[u] & 1u));
			if (lane == 0 && w < words) out[w] = word;
		}
	}
}

void launch_eval_predicate(const PredProgram &prog, long long n, uint32_t *out, long long words, cudaStream_t s,
                           int64_t *launches) {
	if (words <= 0) return;
	const long long blocks = std::min<long long>((words + 8 * PRED_U - 1) / (8 * PRED_U), 148 * 8);
	eval_predicate_kernel<<<(unsigned)blocks, 256, 0, s>>>(prog, n, out, words);
	LB_CUDA(cudaGetLastError());
	if (launches) (*launches)++;
}

// dst[dst_first + i] = src[rows[i]], validity bits set with atomicOr (the destination words start cleared)
__global__ void gather_column_kernel(const void *__restrict__ src, const uint32_t *__restrict__ src_valid,
                                     const int64_t *__restrict__ rows, long long m, int elem_bytes, void *__restrict__ dst,
                                     uint32_t *__restrict__ dst_valid, long long dst_first) {
	const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= m) return;
	const long long r = rows[i], o = dst_first + i;
	if (elem_bytes == 8) static_cast<long long *>(dst)[o] = static_cast<const long long *>(src)[r];
	else static_cast<int32_t *>(dst)[o] = static_cast<const int32_t *>(src)[r];
	if ((src_valid[r >> 5] >> (r & 31)) & 1u) atomicOr(dst_valid + (o >> 5), 1u << (o & 31));
}

void launch_gather_column(const void *src, const uint32_t *src_valid, const int64_t *d_rows, long long m, int elem_bytes,
                          void *dst, uint32_t *dst_valid, long long dst_first, cudaStream_t s, int64_t *launches) {
	if (m <= 0) return;
	gather_column_kernel<<<(unsigned)((m + 255) / 256), 256, 0, s>>>(src, src_valid, d_rows, m, elem_bytes, dst, dst_valid,
	                                                                  dst_first);
	LB_CUDA(cudaGetLastError());
	if (launches) (*launches)++;
}

// remaps dictionary codes after a merge: codes[first + i] = map[codes[first + i]]
__global__ void remap_codes_kernel(int32_t *codes, long long first, long long m, const int32_t *__restrict__ map) {
	const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < m) codes[first + i] = map[codes[first + i]];
}

// ------------------------------------------------------------------------------------------------ storage
DevBitmap::~DevBitmap() {
	if (p) cudaFree(p);
}

ColumnStore::~ColumnStore() {
	for (auto &c : cols_) {
		if (c->data) cudaFree(c->data);
		if (c->valid) cudaFree(c->valid);
	}
}

int ColumnStore::find(const std::string &name) const {
	for (size_t i = 0; i < cols_.size(); i++)
		if (cols_[i]->name == name) return (int)i;
	return -1;
}

void ColumnStore::ensure_cap(Column &c, int64_t cap_rows, cudaStream_t s) {
	cap_rows = (std::max<int64_t>(cap_rows, 256) + 255) & ~255ll;
	if (cap_rows <= c.cap) return;
	void *nd = nullptr;
	uint32_t *nv = nullptr;
	LB_CUDA(cudaMalloc(&nd, (size_t)cap_rows * c.elem_bytes()));
	LB_CUDA(cudaMalloc(reinterpret_cast<void **>(&nv), (size_t)(cap_rows / 32) * sizeof(uint32_t)));
	LB_CUDA(cudaMemsetAsync(nd, 0, (size_t)cap_rows * c.elem_bytes(), s));
	LB_CUDA(cudaMemsetAsync(nv, 0, (size_t)(cap_rows / 32) * sizeof(uint32_t), s));
	if (c.cap > 0) {
		LB_CUDA(cudaMemcpyAsync(nd, c.data, (size_t)c.cap * c.elem_bytes(), cudaMemcpyDeviceToDevice, s));
		LB_CUDA(cudaMemcpyAsync(nv, c.valid, (size_t)(c.cap / 32) * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
	}
	LB_CUDA(cudaStreamSynchronize(s));
	if (c.data) cudaFree(c.data);
	if (c.valid) cudaFree(c.valid);
	c.data = nd;
	c.valid = nv;
	c.cap = cap_rows;
}

ColumnStore::Column &ColumnStore::define(const std::string &name, ColType type) {
	int i = find(name);
	if (i >= 0) {
		if (cols_[i]->type != type) throw Error("extra column \"" + name + "\" exists with another type");
		return *cols_[i];
	}
	if (name.empty()) throw Error("extra column needs a name");
	auto c = std::make_unique<Column>();
	c->name = name;
	c->type = type;
	cols_.push_back(std::move(c));
	epoch_++;
	return *cols_.back();
}

void ColumnStore::define_like(const ColumnStore &other) {
	for (auto &oc : other.cols_) {
		Column &c = define(oc->name, oc->type);
		c.dict = oc->dict;
		c.dict_index = oc->dict_index;
	}
}

void ColumnStore::swap(ColumnStore &other) {
	std::lock_guard<std::mutex> g(cache_mu_);
	cols_.swap(other.cols_);
	cache_.clear();
	other.cache_.clear();
	epoch_ = std::max(epoch_, other.epoch_) + 1;
	other.epoch_ = epoch_;
}

void ColumnStore::grow(int64_t cap_rows, cudaStream_t s) {
	for (auto &c : cols_) ensure_cap(*c, cap_rows, s);
}

// host staging -> device for rows [first_row, first_row + count); `vbytes` one byte per row (1 = valid)
static void upload_rows(ColumnStore::Column &c, int64_t first_row, int64_t count, const void *host_vals,
                        const std::vector<uint8_t> &vbytes, cudaStream_t s) {
	LB_CUDA(cudaMemcpyAsync(static_cast<char *>(c.data) + (size_t)first_row * c.elem_bytes(), host_vals,
	                        (size_t)count * c.elem_bytes(), cudaMemcpyHostToDevice, s));
	// validity: read-modify-write of the words the range touches (the range rarely starts on a word boundary)
	const int64_t w0 = first_row >> 5, w1 = (first_row + count + 31) >> 5;
	std::vector<uint32_t> words((size_t)(w1 - w0));
	LB_CUDA(cudaMemcpyAsync(words.data(), c.valid + w0, words.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
	LB_CUDA(cudaStreamSynchronize(s));
	for (int64_t i = 0; i < count; i++) {
		const int64_t r = first_row + i;
		uint32_t &w = words[(size_t)((r >> 5) - w0)];
		const uint32_t bit = 1u << (r & 31);
		if (vbytes[(size_t)i]) w |= bit;
		else w &= ~bit;
	}
	LB_CUDA(cudaMemcpyAsync(c.valid + w0, words.data(), words.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
	LB_CUDA(cudaStreamSynchronize(s));
}

static int32_t dict_code(ColumnStore::Column &c, const char *p, size_t len) {
	std::string key(p, len);
	auto it = c.dict_index.find(key);
	if (it != c.dict_index.end()) return it->second;
	const int32_t code = (int32_t)c.dict.size();
	c.dict.push_back(key);
	c.dict_index.emplace(std::move(key), code);
	return code;
}

void ColumnStore::write(const std::string &name, ColType type, int64_t first_row, int64_t count, const void *values,
                        const int64_t *offsets, const uint8_t *valid, int64_t cap_rows, cudaStream_t s) {
	if (count <= 0) return;
	if (!values) throw Error("null column values");
	Column &c = define(name, type);
	ensure_cap(c, std::max(cap_rows, first_row + count), s);
	std::vector<uint8_t> vb((size_t)count, 1);
	if (valid)
		for (int64_t i = 0; i < count; i++) vb[(size_t)i] = valid[i] ? 1 : 0;
	if (type == COL_STR) {
		if (!offsets) throw Error("string column needs offsets[count + 1]");
		std::vector<int32_t> codes((size_t)count, 0);
		const char *bytes = static_cast<const char *>(values);
		for (int64_t i = 0; i < count; i++)
			if (vb[(size_t)i]) codes[(size_t)i] = dict_code(c, bytes + offsets[i], (size_t)(offsets[i + 1] - offsets[i]));
		upload_rows(c, first_row, count, codes.data(), vb, s);
	} else if (type == COL_BOOL) {
		std::vector<int64_t> v((size_t)count);
		for (int64_t i = 0; i < count; i++) v[(size_t)i] = static_cast<const uint8_t *>(values)[i] ? 1 : 0;
		upload_rows(c, first_row, count, v.data(), vb, s);
	} else {
		upload_rows(c, first_row, count, values, vb, s);
	}
	epoch_++;
}

void ColumnStore::write_arrow(const std::string &name, ColType type, int64_t first_row, int64_t count, const void *values,
                              int value_width, const void *offsets, bool large_offsets, const uint8_t *valid_bits,
                              int64_t off, int64_t cap_rows, cudaStream_t s) {
	if (count <= 0) return;
	std::vector<uint8_t> vb((size_t)count, 1);
	if (valid_bits)
		for (int64_t i = 0; i < count; i++) vb[(size_t)i] = (valid_bits[(off + i) >> 3] >> ((off + i) & 7)) & 1;
	if (type == COL_STR) {
		// Arrow offsets index the (unsliced) data buffer: rebase them onto a pointer to the first used byte
		std::vector<int64_t> o((size_t)count + 1);
		for (int64_t i = 0; i <= count; i++)
			o[(size_t)i] = large_offsets ? static_cast<const int64_t *>(offsets)[off + i]
			                             : (int64_t) static_cast<const int32_t *>(offsets)[off + i];
		const char *base = values ? static_cast<const char *>(values) : "";
		write(name, type, first_row, count, base, o.data(), vb.data(), cap_rows, s);
		return;
	}
	if (type == COL_BOOL) { // Arrow booleans are bit-packed
		std::vector<uint8_t> v((size_t)count);
		const uint8_t *bits = static_cast<const uint8_t *>(values);
		for (int64_t i = 0; i < count; i++) v[(size_t)i] = (bits[(off + i) >> 3] >> ((off + i) & 7)) & 1;
		write(name, type, first_row, count, v.data(), nullptr, vb.data(), cap_rows, s);
		return;
	}
	// widen to the 8-byte storage type; value_width < 0 = unsigned
	const int wdt = value_width < 0 ? -value_width : value_width;
	const char *base = static_cast<const char *>(values) + (size_t)off * wdt;
	if (type == COL_F64) {
		std::vector<double> v((size_t)count);
		for (int64_t i = 0; i < count; i++)
			v[(size_t)i] = wdt == 4 ? (double)reinterpret_cast<const float *>(base)[i] : reinterpret_cast<const double *>(base)[i];
		write(name, type, first_row, count, v.data(), nullptr, vb.data(), cap_rows, s);
		return;
	}
	std::vector<int64_t> v((size_t)count);
	for (int64_t i = 0; i < count; i++) {
		switch (value_width) {
		case 1: v[(size_t)i] = reinterpret_cast<const int8_t *>(base)[i]; break;
		case -1: v[(size_t)i] = reinterpret_cast<const uint8_t *>(base)[i]; break;
		case 2: v[(size_t)i] = reinterpret_cast<const int16_t *>(base)[i]; break;
		case -2: v[(size_t)i] = reinterpret_cast<const uint16_t *>(base)[i]; break;
		case 4: v[(size_t)i] = reinterpret_cast<const int32_t *>(base)[i]; break;
		case -4: v[(size_t)i] = reinterpret_cast<const uint32_t *>(base)[i]; break;
		default: v[(size_t)i] = reinterpret_cast<const int64_t *>(base)[i]; break;
		}
	}
	write(name, type, first_row, count, v.data(), nullptr, vb.data(), cap_rows, s);
}

// ------------------------------------------------------------------------------------------------ predicate compiler
// Grammar = what ExpressionToLancePredicate emits (src/lance_optimizer.cpp:204-344):
//   col <op> const | const <op> col (=, !=, <>, <, >, <=, >=) | col IS [NOT] NULL | NOT <expr> | col [NOT] IN (c, ..)
//   AND / OR with parentheses; constants as FormatConstant writes them (:138-180).
// Semantics (same as the host mirror's evaluator, csrc/host/predicate.cpp, which the tests use as the checker):
// SQL three-valued logic, a comparison between incomparable types is NULL, only TRUE keeps a row.
namespace {

struct Lit {
	enum Kind { NUL, STR, INT, FLT, BOOL } kind = NUL;
	std::string s;
	int64_t i = 0;
	double d = 0;
};

struct Compiler {
	const std::string &src;
	const ColumnStore &cols;
	size_t pos = 0;
	PredProgram prog;
	std::vector<int64_t> aux_i64;
	std::vector<double> aux_f64;
	std::vector<uint8_t> aux_u8;
	int depth = 0;

	Compiler(const std::string &s, const ColumnStore &c) : src(s), cols(c) {}

	[[noreturn]] void fail(const std::string &why) const {
		throw Error("lance predicate: " + why + " at offset " + std::to_string(pos) + " in \"" + src + "\"");
	}
	void skip() {
		while (pos < src.size() && isspace((unsigned char)src[pos])) pos++;
	}
	bool eof() {
		skip();
		return pos >= src.size();
	}
	bool peek_kw(const char *kw) {
		skip();
		const size_t len = strlen(kw);
		if (pos + len > src.size()) return false;
		for (size_t i = 0; i < len; i++)
			if (toupper((unsigned char)src[pos + i]) != kw[i]) return false;
		return !(pos + len < src.size() && (isalnum((unsigned char)src[pos + len]) || src[pos + len] == '_'));
	}
	bool take_kw(const char *kw) {
		if (!peek_kw(kw)) return false;
		pos += strlen(kw);
		return true;
	}
	bool take_ch(char c) {
		skip();
		if (pos < src.size() && src[pos] == c) {
			pos++;
			return true;
		}
		return false;
	}
	PredOp &emit(int op) {
		if (prog.n_ops >= PRED_MAX_OPS) fail("predicate too long for the device evaluator");
		PredOp &o = prog.ops[prog.n_ops++];
		o = PredOp();
		o.op = op;
		if (op < PRED_NOT) {
			if (++depth > PRED_MAX_DEPTH) fail("predicate nested too deeply");
		} else if (op != PRED_NOT) {
			depth--;
		}
		return o;
	}

	// operand: literal (returns -1) or column (returns its index)
	int operand(Lit &lit) {
		skip();
		if (pos >= src.size()) fail("operand expected");
		const char c = src[pos];
		if (c == '\'') {
			pos++;
			lit.kind = Lit::STR;
			while (true) {
				if (pos >= src.size()) fail("unterminated string");
				if (src[pos] == '\'') {
					if (pos + 1 < src.size() && src[pos + 1] == '\'') {
						lit.s += '\'';
						pos += 2;
						continue;
					}
					pos++;
					break;
				}
				lit.s += src[pos++];
			}
			return -1;
		}
		if (isdigit((unsigned char)c) || c == '-' || c == '+' || c == '.') {
			const size_t b = pos;
			if (c == '-' || c == '+') pos++;
			bool flt = false;
			while (pos < src.size() &&
			       (isdigit((unsigned char)src[pos]) || src[pos] == '.' || src[pos] == 'e' || src[pos] == 'E' ||
			        ((src[pos] == '-' || src[pos] == '+') && (src[pos - 1] == 'e' || src[pos - 1] == 'E')))) {
				if (!isdigit((unsigned char)src[pos])) flt = true;
				pos++;
			}
			const std::string tok = src.substr(b, pos - b);
			if (flt) {
				lit.kind = Lit::FLT;
				lit.d = strtod(tok.c_str(), nullptr);
			} else {
				lit.kind = Lit::INT;
				lit.i = strtoll(tok.c_str(), nullptr, 10);
				lit.d = (double)lit.i;
			}
			return -1;
		}
		if (take_kw("TRUE")) {
			lit.kind = Lit::BOOL, lit.i = 1, lit.d = 1;
			return -1;
		}
		if (take_kw("FALSE")) {
			lit.kind = Lit::BOOL, lit.i = 0, lit.d = 0;
			return -1;
		}
		if (take_kw("NULL")) return -1;
		if (isalpha((unsigned char)c) || c == '_' || c == '"') {
			std::string name;
			if (c == '"') {
				pos++;
				while (pos < src.size() && src[pos] != '"') name += src[pos++];
				if (pos >= src.size()) fail("unterminated identifier");
				pos++;
			} else {
				while (pos < src.size() && (isalnum((unsigned char)src[pos]) || src[pos] == '_')) name += src[pos++];
			}
			const int ci = cols.find(name);
			if (ci < 0) fail("unknown column \"" + name + "\"");
			return ci;
		}
		fail("unexpected character");
	}

	static bool apply(int cmp, int v) { return ((v < 0 ? 1 : (v > 0 ? 4 : 2)) & cmp) != 0; }
	static int flip(int cmp) { return (cmp & 2) | ((cmp & 1) << 2) | ((cmp & 4) >> 2); }
	void bind(PredOp &o, const ColumnStore::Column &c) {
		if (!c.data || !c.valid) fail("column \"" + c.name + "\" holds no rows yet");
		o.data = c.data;
		o.valid = c.valid;
		o.col_is_f64 = c.type == COL_F64;
	}
	void emit_null() { emit(PRED_CONST).ival = 2; }

	void compare_leaf(const ColumnStore::Column &c, int cmp, const Lit &l) {
		if (l.kind == Lit::NUL) return emit_null();
		switch (c.type) {
		case COL_STR: {
			if (l.kind != Lit::STR) return emit_null();
			PredOp &o = emit(PRED_STR_LUT);
			bind(o, c);
			o.aux_off = (int64_t)aux_u8.size();
			for (auto &e : c.dict) {
				const int v = e.compare(l.s);
				aux_u8.push_back(apply(cmp, v < 0 ? -1 : (v > 0 ? 1 : 0)) ? 1 : 0);
			}
			aux_u8.push_back(0); // an empty dictionary still has one addressable entry
			return;
		}
		case COL_I64:
			if (l.kind == Lit::INT || l.kind == Lit::BOOL) {
				PredOp &o = emit(PRED_CMP_I64);
				bind(o, c), o.cmp = cmp, o.ival = l.i;
				return;
			}
			if (l.kind == Lit::FLT) {
				PredOp &o = emit(PRED_CMP_F64);
				bind(o, c), o.cmp = cmp, o.dval = l.d;
				return;
			}
			return emit_null();
		case COL_F64: {
			if (l.kind != Lit::INT && l.kind != Lit::FLT) return emit_null();
			PredOp &o = emit(PRED_CMP_F64);
			bind(o, c), o.cmp = cmp, o.dval = l.d;
			return;
		}
		case COL_BOOL: {
			if (l.kind != Lit::BOOL && l.kind != Lit::INT) return emit_null();
			PredOp &o = emit(PRED_CMP_I64);
			bind(o, c), o.cmp = cmp, o.ival = l.i;
			return;
		}
		}
		emit_null();
	}

	void in_leaf(const ColumnStore::Column &c, bool neg, const std::vector<Lit> &lits) {
		bool saw_null = false;
		for (auto &l : lits) saw_null |= l.kind == Lit::NUL;
		if (c.type == COL_STR) {
			PredOp &o = emit(PRED_STR_LUT);
			bind(o, c);
			o.aux_off = (int64_t)aux_u8.size();
			for (auto &e : c.dict) {
				bool hit = false;
				for (auto &l : lits) hit |= l.kind == Lit::STR && l.s == e;
				aux_u8.push_back(hit ? (neg ? 0 : 1) : (saw_null ? 2 : (neg ? 1 : 0)));
			}
			aux_u8.push_back(0);
			return;
		}
		// comparable literals only (the others never match: the host mirror's compare() says "not comparable")
		bool any_flt = c.type == COL_F64;
		std::vector<const Lit *> use;
		for (auto &l : lits) {
			const bool ok = c.type == COL_I64   ? (l.kind == Lit::INT || l.kind == Lit::BOOL || l.kind == Lit::FLT)
			                : c.type == COL_F64 ? (l.kind == Lit::INT || l.kind == Lit::FLT)
			                                    : (l.kind == Lit::BOOL || l.kind == Lit::INT);
			if (!ok) continue;
			any_flt |= l.kind == Lit::FLT;
			use.push_back(&l);
		}
		PredOp &o = emit(any_flt ? PRED_IN_F64 : PRED_IN_I64);
		bind(o, c);
		o.negate = neg, o.has_null = saw_null, o.n_list = (int32_t)use.size();
		if (any_flt) {
			o.aux_off = (int64_t)aux_f64.size();
			for (auto *l : use) aux_f64.push_back(l->d);
		} else {
			o.aux_off = (int64_t)aux_i64.size();
			for (auto *l : use) aux_i64.push_back(l->i);
		}
	}

	void primary() {
		if (take_ch('(')) {
			expr();
			if (!take_ch(')')) fail("')' expected");
			return;
		}
		Lit la;
		const int ca = operand(la);
		if (ca >= 0 && peek_kw("IS")) {
			take_kw("IS");
			const bool neg = take_kw("NOT");
			if (!take_kw("NULL")) fail("NULL expected after IS");
			bind(emit(neg ? PRED_NOT_NULL : PRED_IS_NULL), cols.column((size_t)ca));
			return;
		}
		if (ca >= 0 && (peek_kw("IN") || peek_kw("NOT"))) {
			const bool neg = take_kw("NOT");
			if (!take_kw("IN")) fail("IN expected");
			if (!take_ch('(')) fail("'(' expected after IN");
			std::vector<Lit> lits;
			do {
				Lit l;
				if (operand(l) >= 0) fail("constant expected in IN list");
				lits.push_back(l);
			} while (take_ch(','));
			if (!take_ch(')')) fail("')' expected");
			in_leaf(cols.column((size_t)ca), neg, lits);
			return;
		}
		skip();
		static const struct {
			const char *s;
			int cmp;
		} ops[] = {{"<=", CMP_LE}, {">=", CMP_GE}, {"!=", CMP_NE}, {"<>", CMP_NE}, {"=", CMP_EQ}, {"<", CMP_LT}, {">", CMP_GT}};
		int cmp = -1;
		for (auto &o : ops)
			if (src.compare(pos, strlen(o.s), o.s) == 0) {
				cmp = o.cmp;
				pos += strlen(o.s);
				break;
			}
		if (cmp < 0) fail("comparison operator expected");
		Lit lb;
		const int cb = operand(lb);
		if ((ca >= 0) == (cb >= 0)) fail("a comparison needs one column and one constant"); // :216-232
		if (ca >= 0) compare_leaf(cols.column((size_t)ca), cmp, lb);
		else compare_leaf(cols.column((size_t)cb), flip(cmp), la);
	}
	void not_expr() {
		if (take_kw("NOT")) {
			not_expr();
			emit(PRED_NOT);
			return;
		}
		primary();
	}
	void and_expr() {
		not_expr();
		while (take_kw("AND")) {
			not_expr();
			emit(PRED_AND);
		}
	}
	void expr() {
		and_expr();
		while (take_kw("OR")) {
			and_expr();
			emit(PRED_OR);
		}
	}
};

struct TmpDev {
	void *p = nullptr;
	~TmpDev() {
		if (p) cudaFree(p);
	}
	void upload(const void *h, size_t bytes, cudaStream_t s) {
		if (!bytes) return;
		LB_CUDA(cudaMalloc(&p, bytes));
		LB_CUDA(cudaMemcpyAsync(p, h, bytes, cudaMemcpyHostToDevice, s));
	}
};

} // namespace

std::shared_ptr<DevBitmap> ColumnStore::bitmap(const std::string &predicate, int64_t n_rows, cudaStream_t s,
                                               int64_t *launches, bool *was_cached) {
	std::lock_guard<std::mutex> g(cache_mu_);
	auto it = cache_.find(predicate);
	if (it != cache_.end() && it->second.epoch == epoch_ && it->second.n_rows == n_rows) {
		it->second.last_use = ++use_clock_;
		if (was_cached) *was_cached = true;
		return it->second.bm;
	}
	if (was_cached) *was_cached = false;
	Compiler c(predicate, *this);
	c.expr();
	if (!c.eof()) c.fail("trailing input");
	TmpDev a64, af, a8;
	a64.upload(c.aux_i64.data(), c.aux_i64.size() * sizeof(int64_t), s);
	af.upload(c.aux_f64.data(), c.aux_f64.size() * sizeof(double), s);
	a8.upload(c.aux_u8.data(), c.aux_u8.size(), s);
	c.prog.aux_i64 = static_cast<const int64_t *>(a64.p);
	c.prog.aux_f64 = static_cast<const double *>(af.p);
	c.prog.aux_u8 = static_cast<const uint8_t *>(a8.p);
	auto bm = std::make_shared<DevBitmap>();
	bm->bits = n_rows;
	bm->words = (n_rows + 31) / 32;
	LB_CUDA(cudaMalloc(&bm->p, (size_t)std::max<int64_t>(bm->words, 1) * sizeof(uint32_t)));
	launch_eval_predicate(c.prog, n_rows, static_cast<uint32_t *>(bm->p), bm->words, s, launches);
	LB_CUDA(cudaStreamSynchronize(s)); // the aux tables die with this frame; the bitmap is then usable from any stream
	if (cache_.size() >= 16) { // small LRU: a workload's WHERE clauses are few and constant
		auto victim = cache_.begin();
		for (auto j = cache_.begin(); j != cache_.end(); ++j)
			if (j->second.last_use < victim->second.last_use) victim = j;
		cache_.erase(victim);
	}
	Cached e;
	e.epoch = epoch_, e.n_rows = n_rows, e.bm = bm, e.last_use = ++use_clock_;
	cache_[predicate] = e;
	return bm;
}

// ------------------------------------------------------------------------------------------------ persistence, merge
static const char COL_MAGIC[8] = {'L', 'B', 'C', 'O', 'L', 'S', '1', '\0'};

void ColumnStore::save(FILE *f, int64_t n_rows, cudaStream_t s) const {
	int64_t nc = (int64_t)cols_.size();
	bool ok = fwrite(COL_MAGIC, 1, 8, f) == 8 && fwrite(&nc, sizeof(nc), 1, f) == 1;
	const size_t words = (size_t)((n_rows + 31) / 32);
	for (auto &c : cols_) {
		int32_t h[2] = {(int32_t)c->type, (int32_t)c->name.size()};
		ok = ok && fwrite(h, sizeof(int32_t), 2, f) == 2 && fwrite(c->name.data(), 1, c->name.size(), f) == c->name.size();
		int64_t nd = (int64_t)c->dict.size();
		ok = ok && fwrite(&nd, sizeof(nd), 1, f) == 1;
		for (auto &e : c->dict) {
			int32_t len = (int32_t)e.size();
			ok = ok && fwrite(&len, sizeof(len), 1, f) == 1 && fwrite(e.data(), 1, e.size(), f) == e.size();
		}
		std::vector<char> buf((size_t)n_rows * c->elem_bytes());
		std::vector<uint32_t> vb(words);
		if (n_rows > 0 && c->cap >= n_rows) {
			LB_CUDA(cudaMemcpyAsync(buf.data(), c->data, buf.size(), cudaMemcpyDeviceToHost, s));
			LB_CUDA(cudaMemcpyAsync(vb.data(), c->valid, words * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
			LB_CUDA(cudaStreamSynchronize(s));
		}
		ok = ok && fwrite(buf.data(), 1, buf.size(), f) == buf.size() && fwrite(vb.data(), sizeof(uint32_t), words, f) == words;
	}
	if (!ok) throw Error("short write of the extra columns");
}

void ColumnStore::load(FILE *f, const std::string &path, int64_t n_rows, int64_t cap_rows, cudaStream_t s) {
	char magic[8];
	int64_t nc = 0;
	if (fread(magic, 1, 8, f) != 8 || memcmp(magic, COL_MAGIC, 8) != 0 || fread(&nc, sizeof(nc), 1, f) != 1 || nc < 0 ||
	    nc > 4096)
		throw Error("corrupt extra-column section: " + path);
	const size_t words = (size_t)((n_rows + 31) / 32);
	for (int64_t i = 0; i < nc; i++) {
		int32_t h[2];
		if (fread(h, sizeof(int32_t), 2, f) != 2 || h[0] < 0 || h[0] > COL_STR || h[1] <= 0 || h[1] > 4096)
			throw Error("corrupt extra-column header: " + path);
		std::string name((size_t)h[1], '\0');
		int64_t nd = 0;
		if (fread(&name[0], 1, name.size(), f) != name.size() || fread(&nd, sizeof(nd), 1, f) != 1 || nd < 0 || nd > n_rows)
			throw Error("truncated extra column: " + path);
		Column &c = define(name, (ColType)h[0]);
		for (int64_t j = 0; j < nd; j++) {
			int32_t len = 0;
			if (fread(&len, sizeof(len), 1, f) != 1 || len < 0 || len > (1 << 28)) throw Error("truncated extra column: " + path);
			std::string e((size_t)len, '\0');
			if (len && fread(&e[0], 1, e.size(), f) != e.size()) throw Error("truncated extra column: " + path);
			c.dict_index.emplace(e, (int32_t)c.dict.size());
			c.dict.push_back(std::move(e));
		}
		ensure_cap(c, std::max<int64_t>(cap_rows, n_rows), s);
		std::vector<char> buf((size_t)n_rows * c.elem_bytes());
		std::vector<uint32_t> vb(words);
		if (fread(buf.data(), 1, buf.size(), f) != buf.size() || fread(vb.data(), sizeof(uint32_t), words, f) != words)
			throw Error("truncated extra column: " + path);
		if (c.type == COL_STR) // a code outside the dictionary would index past the predicate's lookup table
			for (int64_t r = 0; r < n_rows; r++) {
				const int32_t code = reinterpret_cast<const int32_t *>(buf.data())[r];
				if (code < 0 || code >= std::max<int64_t>(nd, 1)) throw Error("corrupt string codes: " + path);
			}
		if (n_rows > 0) {
			LB_CUDA(cudaMemcpyAsync(c.data, buf.data(), buf.size(), cudaMemcpyHostToDevice, s));
			LB_CUDA(cudaMemcpyAsync(c.valid, vb.data(), words * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
			LB_CUDA(cudaStreamSynchronize(s));
		}
	}
	epoch_++;
}

void ColumnStore::gather_from(const ColumnStore &src, const int64_t *d_rows, int64_t m, int64_t dst_first, int64_t cap_rows,
                              cudaStream_t s, int64_t *launches) {
	if (m <= 0) return;
	for (auto &sc : src.cols_) {
		if (!sc->data) continue;
		Column &c = define(sc->name, sc->type);
		ensure_cap(c, std::max(cap_rows, dst_first + m), s);
		launch_gather_column(sc->data, sc->valid, d_rows, m, c.elem_bytes(), c.data, c.valid, dst_first, s, launches);
		if (c.type == COL_STR && !sc->dict.empty()) { // source codes -> this column's dictionary
			std::vector<int32_t> map(sc->dict.size());
			for (size_t j = 0; j < sc->dict.size(); j++) map[j] = dict_code(c, sc->dict[j].data(), sc->dict[j].size());
			TmpDev dm;
			dm.upload(map.data(), map.size() * sizeof(int32_t), s);
			remap_codes_kernel<<<(unsigned)((m + 255) / 256), 256, 0, s>>>(static_cast<int32_t *>(c.data), dst_first, m,
			                                                               static_cast<const int32_t *>(dm.p));
			LB_CUDA(cudaGetLastError());
			if (launches) (*launches)++;
			LB_CUDA(cudaStreamSynchronize(s));
		}
	}
	LB_CUDA(cudaStreamSynchronize(s));
	epoch_++;
}

} // namespace lb
