// cudasp_extension.cpp -- the DuckDB side of the drop-in: table function `cudasp_scan` over the C ABI.
//
// Plain C++ (no nvcc): everything GPU-related lives behind include/cudasp.h (libcudasp_b200.so).
// Same extension name (`cudasp`), entry point (`cudasp_duckdb_cpp_init`), function name, positional and
// named parameters, input columns, output schema, argument checks and error texts as the reference's
// src/cudasp_extension.cu (bind :420-509, states :87-152, accumulate :154-224, in-out protocol :534-624,
// registration :626-659).  What differs is the data path: rows are written straight from the DuckDB
// vectors into a pinned staging slot of the C ABI (no per-row std::string, no limb conversion, no
// transposition), a full slot is submitted asynchronously, and results (row ids) are collected whenever a
// worker has nothing better to do.  GPU errors THROW (the reference printf()s and drops the batch, :325,378).
#include "cudasp_extension.hpp"

#include "duckdb.hpp"
#include "duckdb/common/exception.hpp"
#include "duckdb/function/table_function.hpp"
#include "duckdb/main/extension/extension_loader.hpp"
#include "duckdb/parallel/task_scheduler.hpp"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "../../include/cudasp.h"
#include "../../include/cudasp_parquet.h"

#if defined(__SSE2__) && !defined(CUDASP_NO_STREAMING_STORES)
#include <emmintrin.h>
#define CUDASP_NT 1
#else
#define CUDASP_NT 0
#endif

namespace duckdb {

namespace {

constexpr idx_t SCALAR_SIZE = 32;      // scan_private_key: 32 B little-endian
constexpr idx_t POINT_SIZE = 64;       // x_LE || y_LE
constexpr idx_t TXID_SIZE = 32;        // the usual txid length (any length is accepted)
constexpr idx_t SEC1_SIZE = 33;        // (02|03) || x_BE: the index's own storage format, decompressed on the GPU
constexpr idx_t DEFAULT_BATCH = 300000;  // reference default (:473)
constexpr idx_t MAX_BATCH = 10000000;

struct CudaspScanBindData : public TableFunctionData {
	string scan_private_key;
	string spend_public_key;
	string label_keys;  // n_labels * 64 B, concatenated
	uint32_t n_labels = 0;
	idx_t batch_size = DEFAULT_BATCH;
	bool emit_early = false;  // emit_early := true (extension): matches leave during Execute, not only in Final
	bool details = false;  // details := true (extension): one result row per found output, + k, output_index, label_index, t_k
};

// The staged keys (pinned slot) and the kept txids are written once and not read by this CPU again -- the GPU's copy engine reads the
// slot, a txid is read back only for the handful of rows that match -- so they are written with streaming (non-temporal) stores: no
// read-for-ownership of the destination line, no cache pollution.  With 32 workers feeding 8 GPUs the extension moves ~130 B per row
// each way, and host memory bandwidth is what the workers share.  dst must be 16-byte aligned; FenceStreamingStores() before the
// data is handed on.
template <idx_t BYTES>
static inline void CopyStreaming(void *dst, const void *src) {
	static_assert(BYTES % 16 == 0, "whole 16-byte blocks");
#if CUDASP_NT
	for (idx_t i = 0; i < BYTES / 16; i++) {
		_mm_stream_si128(reinterpret_cast<__m128i *>(dst) + i, _mm_loadu_si128(reinterpret_cast<const __m128i *>(src) + i));
	}
#else
	memcpy(dst, src, BYTES);
#endif
}
static inline void FenceStreamingStores() {
#if CUDASP_NT
	_mm_sfence();
#endif
}

// Growable byte buffer that does NOT value-initialise what it grows by (vector::resize would write every txid byte twice)
// and keeps its capacity when a batch's metadata is recycled.
template <class T>
struct RawBuffer {
	T *ptr = nullptr;
	size_t len = 0, cap = 0;
	RawBuffer() = default;
	RawBuffer(const RawBuffer &) = delete;
	RawBuffer &operator=(const RawBuffer &) = delete;
	~RawBuffer() {
		free(ptr);
	}
	void Reserve(size_t n) {
		if (n > cap) {
			const size_t want = MaxValue<size_t>(n, cap + cap / 2);
			T *p = static_cast<T *>(realloc(ptr, want * sizeof(T)));
			if (!p) {
				throw std::bad_alloc();
			}
			ptr = p;
			cap = want;
		}
	}
	// make room for n more elements and return where they go
	T *Grow(size_t n) {
		Reserve(len + n);
		T *at = ptr + len;
		len += n;
		return at;
	}
	void Truncate(size_t n) {
		len = n;
	}
};

// Host-side columns of one submitted batch that never go to the GPU: only needed for matched rows.
struct InFlightBatch {
	// txid i = txid_bytes[txid_end[i-1] .. txid_end[i]); while every txid so far is TXID_SIZE bytes (the normal case) the
	// index is implicit (txid i at TXID_SIZE * i) and txid_end stays empty
	RawBuffer<char> txid_bytes;
	vector<uint32_t> txid_end;
	bool txid_fixed = true;
	RawBuffer<int32_t> heights;
	uint32_t rows = 0;

	void Clear() {
		txid_bytes.Truncate(0);
		txid_end.clear();
		heights.Truncate(0);
		txid_fixed = true;
		rows = 0;
	}
	// leave the implicit index: write the explicit one for the rows added so far
	void MakeTxidIndexExplicit(idx_t txid_size) {
		if (!txid_fixed) {
			return;
		}
		txid_fixed = false;
		txid_end.resize(rows);
		for (uint32_t i = 0; i < rows; i++) {
			txid_end[i] = (uint32_t)((i + 1) * txid_size);
		}
	}
	void AddTxid(const string_t &tx, idx_t txid_size) {
		const idx_t n = tx.GetSize();
		if (txid_fixed && n != txid_size) {
			MakeTxidIndexExplicit(txid_size);
		}
		memcpy(txid_bytes.Grow(n), tx.GetData(), n);
		if (!txid_fixed) {
			txid_end.push_back((uint32_t)txid_bytes.len);
		}
	}
	const uint8_t *tweak_keys = nullptr;  // the slot's pinned buffer: valid until the result is released
	uint32_t key_bytes = 64;              // one key format per batch (CUDASP_KEYS_XY64 / CUDASP_KEYS_SEC1_33)
};

struct CudaspScanState : public GlobalTableFunctionState {
	~CudaspScanState() override {
		if (query) {
			cudasp_query_destroy(query);
		}
	}
	idx_t MaxThreads() const override {
		return GlobalTableFunctionState::MAX_THREADS;
	}

	cudasp_query *query = nullptr;
	std::atomic<uint64_t> next_tag {1};
	std::mutex inflight_lock;
	std::unordered_map<uint64_t, unique_ptr<InFlightBatch>> inflight;
	vector<unique_ptr<InFlightBatch>> meta_pool;  // collected batches' host columns, recycled (their vectors keep capacity)
	// Every worker's first batches are cut short (1/8, 1/4, 1/2 of the steady size, see EnsureSlot) so that the GPUs start
	// while the workers are still filling their first full batch; batch_size stays the maximum.
	std::atomic<uint32_t> batches_started {0};  // statistics only
	uint32_t n_devices = 1;
	uint32_t batch_size = 300000;
	// Every worker fills its own batch, so threads x batch / 2 rows are still on the host when the input ends and the
	// GPUs process them as a tail.  With many workers per GPU the batches are therefore cut to
	// batch_size * 2 * devices / threads rows, but not below one full wave of K1 (148 SMs x 4 blocks x 128 rows on B200):
	// batches of different workers run side by side on their slots' streams.
	uint32_t steady_limit = 300000;

	// protocol state, as in the reference: nothing is emitted before every worker has finalized (:554-591)
	std::atomic<uint64_t> currently_adding {0};
	std::mutex finalize_lock;
	std::atomic<bool> output_thread_claimed {false};

	std::mutex output_lock;
	vector<string> output_txids;
	vector<int32_t> output_heights;
	vector<string> output_tweak_keys;
	bool details = false;
	vector<cudasp_match_detail> output_details;  // details := true: parallel to the three vectors above
	idx_t output_position = 0;
};

struct CudaspScanLocalState : public LocalTableFunctionState {
	bool finalized = false;
	bool is_output_thread = false;
	bool have_slot = false;
	bool input_consumed = false;  // emit_early: the current input chunk is already staged (we are being re-called for output)
	double accumulate_s = 0, first_input = 0;
	double slot_wait_s = 0;  // trace: time EnsureSlot spent waiting for the GPUs (every slot busy)
	uint64_t rows_seen = 0, batches = 0;
	uint32_t chunks_since_poll = 0;
	uint64_t rows_fast = 0;  // rows staged by AccumulateInput's column-wise fast path (trace)
	cudasp_slot slot;
	uint32_t rows = 0;
	uint64_t n_outputs = 0;
	uint32_t key_bytes = 64;   // key format of the batch being filled
	uint32_t batch_limit = 0;  // rows after which the batch being filled is submitted (<= batch_size, see EnsureSlot)
	unique_ptr<InFlightBatch> meta;
};

// CUDASP_TRACE=1 prints coarse host-side timings to stderr (query setup, per-worker accumulate time, final drain)
bool TraceOn() {
	static const bool on = getenv("CUDASP_TRACE") != nullptr;
	return on;
}
double NowSec() {
	return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

[[noreturn]] void ThrowGpu(CudaspScanState &state, const char *what) {
	const char *msg = cudasp_last_error(state.query);
	throw IOException("cudasp_scan: %s failed: %s", what, msg ? msg : "unknown error");
}

}  // namespace

// ------------------------------------------------------------------------------------------ bind
// The key arguments (positions 1..3) and batch_size, shared by cudasp_scan and cudasp_scan_parquet: same checks, same order, same
// messages as the reference's CudaspScanBind (:423-487).
static void BindKeys(TableFunctionBindInput &input, CudaspScanBindData &result) {
	auto &sk = input.inputs[1];
	if (sk.type().id() != LogicalTypeId::BLOB) {
		throw InvalidInputException("Second argument must be a BLOB (32-byte scan_private_key)");
	}
	result.scan_private_key = StringValue::Get(sk);
	if (result.scan_private_key.size() != SCALAR_SIZE) {
		throw InvalidInputException("scan_private_key must be exactly 32 bytes, got %llu bytes",
		                            (uint64_t)result.scan_private_key.size());
	}

	auto &spend = input.inputs[2];
	if (spend.type().id() != LogicalTypeId::BLOB) {
		throw InvalidInputException("Third argument must be a BLOB (64-byte spend_public_key)");
	}
	result.spend_public_key = StringValue::Get(spend);
	if (result.spend_public_key.size() != POINT_SIZE) {
		throw InvalidInputException("spend_public_key must be exactly 64 bytes, got %llu bytes",
		                            (uint64_t)result.spend_public_key.size());
	}

	auto &labels = input.inputs[3];
	if (labels.type().id() != LogicalTypeId::LIST) {
		throw InvalidInputException("Fourth argument must be a LIST[BLOB] (label keys)");
	}
	if (!labels.IsNull()) {
		for (auto &child : ListValue::GetChildren(labels)) {
			if (child.type().id() != LogicalTypeId::BLOB) {
				throw InvalidInputException("All elements in label_keys must be BLOBs");
			}
			const string &key = StringValue::Get(child);
			if (key.size() != POINT_SIZE) {
				throw InvalidInputException("Each label key must be exactly 64 bytes, got %llu bytes", (uint64_t)key.size());
			}
			result.label_keys += key;
			result.n_labels++;
		}
	}
	if (result.n_labels > CUDASP_MAX_LABELS) {
		throw InvalidInputException("cudasp_scan supports at most %llu label keys", (uint64_t)CUDASP_MAX_LABELS);
	}

	auto bs = input.named_parameters.find("batch_size");
	if (bs != input.named_parameters.end()) {
		auto tid = bs->second.type().id();
		if (tid != LogicalTypeId::INTEGER && tid != LogicalTypeId::BIGINT) {
			throw InvalidInputException("batch_size parameter must be an INTEGER");
		}
		int64_t v = bs->second.GetValue<int64_t>();
		if (v <= 0) {
			throw InvalidInputException("batch_size must be positive, got %lld", (long long)v);
		}
		if (v > (int64_t)MAX_BATCH) {
			throw InvalidInputException("batch_size too large (max 10,000,000), got %lld", (long long)v);
		}
		result.batch_size = (idx_t)v;
	}

}

// Same checks, same order, same messages as the reference's CudaspScanBind (:420-509).
static unique_ptr<FunctionData> CudaspScanBind(ClientContext &context, TableFunctionBindInput &input,
                                               vector<LogicalType> &return_types, vector<string> &names) {
	if (input.inputs.size() != 4) {
		throw InvalidInputException("cudasp_scan requires 4 arguments: TABLE, scan_private_key BLOB, spend_public_key "
		                            "BLOB, and label_keys LIST[BLOB]");
	}
	auto result = make_uniq<CudaspScanBindData>();
	BindKeys(input, *result);

	auto ee = input.named_parameters.find("emit_early");
	if (ee != input.named_parameters.end()) {
		if (ee->second.type().id() != LogicalTypeId::BOOLEAN) {
			throw InvalidInputException("emit_early parameter must be a BOOLEAN");
		}
		result->emit_early = !ee->second.IsNull() && ee->second.GetValue<bool>();
	}
	auto dt = input.named_parameters.find("details");
	if (dt != input.named_parameters.end()) {
		if (dt->second.type().id() != LogicalTypeId::BOOLEAN) {
			throw InvalidInputException("details parameter must be a BOOLEAN");
		}
		result->details = !dt->second.IsNull() && dt->second.GetValue<bool>();
	}

	return_types = {LogicalType::BLOB, LogicalType::INTEGER, LogicalType::BLOB};
	names = {"txid", "height", "tweak_key"};
	if (result->details) {
		// not in the reference: the BIP-352 k-loop on matching rows (include/cudasp.h, cudasp_match_detail)
		return_types.insert(return_types.end(), {LogicalType::UINTEGER, LogicalType::UINTEGER, LogicalType::INTEGER, LogicalType::BLOB});
		names.insert(names.end(), {"k", "output_index", "label_index", "t_k"});
	}
	return std::move(result);
}

// ------------------------------------------------------------------------------------------ init
// Global init does the once-per-query work the reference repeats per batch (key conversion, :277-297) and per
// process (Solver::initialize, gsv.h:1094-1122): cudasp_query_create recodes the scan key, uploads tables,
// allocates pinned staging + device buffers on every visible GPU.
static unique_ptr<GlobalTableFunctionState> CudaspScanInit(ClientContext &context, TableFunctionInitInput &input) {
	auto &bind = input.bind_data->Cast<CudaspScanBindData>();
	auto state = make_uniq<CudaspScanState>();
	int n_dev = cudasp_device_count();
	if (n_dev <= 0) {
		throw IOException("cudasp_scan: no CUDA device available (this extension has no CPU path)");
	}
	idx_t threads = MaxValue<idx_t>(1, NumericCast<idx_t>(TaskScheduler::GetScheduler(context).NumberOfThreads()));
	// rows per GPU batch: batch_size, cut down when there are many workers per GPU (see steady_limit)
	{
		// one full wave of the ladder kernel: SMs x 4 resident blocks x 128 rows (148 SMs on B200)
		const int sms = cudasp_device_sm_count(0);
		const uint64_t wave = (uint64_t)(sms > 0 ? sms : 148) * 4 * 128;
		const uint64_t share = (uint64_t)bind.batch_size * 2 * (uint64_t)n_dev / threads;
		state->steady_limit = (uint32_t)MinValue<uint64_t>(bind.batch_size, MaxValue<uint64_t>(wave, share));
	}
	// every worker can hold one slot while filling it; two more per device keep a copy and a batch in flight
	uint32_t slots = (uint32_t)MinValue<idx_t>(64, (threads + n_dev - 1) / n_dev + 2);
	// cap the pinned/device footprint for very large batch sizes (~200 B of staging per row and slot)
	while (slots > 3 && (uint64_t)slots * state->steady_limit > 48000000ull) {
		slots--;
	}
	const double t_create = NowSec();
	// the slots are sized for the batches that will actually be submitted (steady_limit rows, room for 8 outputs per row
	// on average before a batch is cut early): less pinned and device memory to allocate on a query's first batches
	int rc = cudasp_query_create_ex(const_data_ptr_cast(bind.scan_private_key.data()),
	                                const_data_ptr_cast(bind.spend_public_key.data()),
	                                bind.n_labels ? const_data_ptr_cast(bind.label_keys.data()) : nullptr, bind.n_labels,
	                                nullptr, 0, state->steady_limit, 8ull * state->steady_limit, slots, &state->query);
	if (rc != CUDASP_OK) {
		const char *msg = cudasp_last_error(nullptr);
		if (rc == CUDASP_ERR_KEY || rc == CUDASP_ERR_ARG) {
			throw InvalidInputException("cudasp_scan: %s", msg ? msg : "invalid keys");
		}
		throw IOException("cudasp_scan: GPU initialisation failed: %s", msg ? msg : "unknown error");
	}
	state->details = bind.details;
	state->n_devices = (uint32_t)n_dev;
	state->batch_size = (uint32_t)bind.batch_size;
	if (bind.details && cudasp_query_enable_details(state->query, 0) != CUDASP_OK) {
		ThrowGpu(*state, "cudasp_query_enable_details");
	}
	if (TraceOn()) {
		fprintf(stderr, "[cudasp] query_create: %.3f s (%u slots/device, %d devices, %llu threads)\n", NowSec() - t_create, slots,
		        n_dev, (unsigned long long)threads);
	}
	return std::move(state);
}

static unique_ptr<LocalTableFunctionState> CudaspScanLocalInit(ExecutionContext &context, TableFunctionInitInput &input,
                                                               GlobalTableFunctionState *global_state) {
	auto &state = global_state->Cast<CudaspScanState>();
	state.currently_adding++;
	return make_uniq<CudaspScanLocalState>();
}

// ------------------------------------------------------------------------------------------ results
// Copy the matched rows of one finished batch into the global result vectors, then recycle its slot.
// Replaces the O(batch) scan of managed match flags under the output mutex (:383-391).
static void CollectResult(CudaspScanState &state, const cudasp_result &r) {
	unique_ptr<InFlightBatch> meta;
	{
		std::lock_guard<std::mutex> guard(state.inflight_lock);
		auto it = state.inflight.find(r.tag);
		if (it == state.inflight.end()) {
			throw InternalException("cudasp_scan: result for an unknown batch tag");
		}
		meta = std::move(it->second);
		state.inflight.erase(it);
	}
	const cudasp_match_detail *det = nullptr;
	uint32_t n_det = 0;
	if (state.details && cudasp_result_details(state.query, &r, &det, &n_det) != CUDASP_OK) {
		ThrowGpu(state, "cudasp_result_details");
	}
	const uint32_t n_rows_out = state.details ? n_det : r.n_matches;
	if (n_rows_out) {
		std::lock_guard<std::mutex> guard(state.output_lock);
		for (uint32_t k = 0; k < n_rows_out; k++) {
			const uint32_t i = state.details ? det[k].row : r.match_ids[k];
			const uint32_t b = meta->txid_fixed ? (uint32_t)(i * TXID_SIZE) : (i ? meta->txid_end[i - 1] : 0);
			const uint32_t e = meta->txid_fixed ? (uint32_t)((i + 1) * TXID_SIZE) : meta->txid_end[i];
			state.output_txids.emplace_back(meta->txid_bytes.ptr + b, e - b);
			state.output_heights.push_back(meta->heights.ptr[i]);
			state.output_tweak_keys.emplace_back(const_char_ptr_cast(meta->tweak_keys) + (size_t)i * meta->key_bytes, meta->key_bytes);
			if (state.details) {
				state.output_details.push_back(det[k]);
			}
		}
	}
	if (cudasp_release_result(state.query, &r) != CUDASP_OK) {
		ThrowGpu(state, "cudasp_release_result");
	}
	meta->Clear();
	std::lock_guard<std::mutex> guard(state.inflight_lock);
	state.meta_pool.push_back(std::move(meta));
}

// Collect every batch that has already finished (block == false) or every batch in flight (block == true).
static void Drain(CudaspScanState &state, bool block) {
	cudasp_result r;
	for (;;) {
		int rc = block ? cudasp_wait(state.query, &r) : cudasp_poll(state.query, &r);
		if (rc < 0) {
			ThrowGpu(state, block ? "cudasp_wait" : "cudasp_poll");
		}
		if (rc == 0) {
			return;
		}
		CollectResult(state, r);
	}
}

static void EnsureSlot(CudaspScanState &state, CudaspScanLocalState &local) {
	if (local.have_slot) {
		return;
	}
	for (;;) {
		int rc = cudasp_try_acquire_slot(state.query, &local.slot);
		if (rc < 0) {
			ThrowGpu(state, "cudasp_try_acquire_slot");
		}
		if (rc == 1) {
			break;
		}
		// every slot is busy: help by collecting the oldest batch in flight (this frees a slot); if nothing is in
		// flight the slots are all being filled by other workers -- block until one of them is recycled.
		const double t_wait = TraceOn() ? NowSec() : 0;
		struct WaitTimer {
			CudaspScanLocalState &l;
			double t0;
			~WaitTimer() {
				if (TraceOn()) {
					l.slot_wait_s += NowSec() - t0;
				}
			}
		} wait_timer {local, t_wait};
		cudasp_result r;
		rc = cudasp_wait(state.query, &r);
		if (rc < 0) {
			ThrowGpu(state, "cudasp_wait");
		}
		if (rc == 1) {
			CollectResult(state, r);
		} else {
			if (cudasp_acquire_slot(state.query, &local.slot) != CUDASP_OK) {
				ThrowGpu(state, "cudasp_acquire_slot");
			}
			break;
		}
	}
	local.have_slot = true;
	local.rows = 0;
	local.n_outputs = 0;
	local.slot.out_offsets[0] = 0;
	{
		std::lock_guard<std::mutex> guard(state.inflight_lock);
		if (!state.meta_pool.empty()) {
			local.meta = std::move(state.meta_pool.back());
			state.meta_pool.pop_back();
		}
	}
	if (!local.meta) {
		local.meta = make_uniq<InFlightBatch>();
	}
	// Every worker ramps its OWN batches up: 1/8, 1/4, 1/2 of the steady size (not below 16 K rows), then full.  The GPUs start
	// after a few milliseconds of input instead of after a worker's first full batch, and a table that is small against
	// threads x batch (config C2: 1 M rows on 16 workers) is mostly scanned by the time its last row has been decoded, instead
	// of arriving as one partial batch per worker in Final.  Small batches of different workers run side by side on their
	// slots' streams.  (Round 1 ramped only the first batches of the whole query: one round of devices per size.)
	const uint32_t k = (uint32_t)MinValue<uint64_t>(local.batches++, 3);
	local.batch_limit = MaxValue<uint32_t>(MinValue<uint32_t>(state.steady_limit, 16384u), state.steady_limit >> (3 - k));
	state.batches_started++;
	local.meta->txid_bytes.Reserve((size_t)local.batch_limit * TXID_SIZE);
	local.meta->heights.Reserve(local.batch_limit);
	local.meta->tweak_keys = local.slot.tweak_keys;
}

// Hand the filled part of the slot to the GPU (async H2D + kernels + D2H on the slot's stream).  Replaces ProcessBatch
// (:226-410): no marshalling is left to do, the slot already holds the device layout.
static void SubmitSlot(CudaspScanState &state, CudaspScanLocalState &local) {
	if (!local.have_slot || local.rows == 0) {
		return;
	}
	const uint64_t tag = state.next_tag++;
	{
		std::lock_guard<std::mutex> guard(state.inflight_lock);
		state.inflight[tag] = std::move(local.meta);
	}
	FenceStreamingStores();
	const int fmt = local.key_bytes == SEC1_SIZE ? CUDASP_KEYS_SEC1_33 : CUDASP_KEYS_XY64;
	if (cudasp_submit_slot_fmt(state.query, &local.slot, local.rows, local.n_outputs, tag, fmt) != CUDASP_OK) {
		ThrowGpu(state, "cudasp_submit_slot");
	}
	local.have_slot = false;
	local.rows = 0;
	local.n_outputs = 0;
}

// ------------------------------------------------------------------------------------------ execute
// One DataChunk (<= 2048 rows) -> pinned staging.  Row rules are the reference's AccumulateInput (:154-224):
// a row with NULL txid, height or tweak_key is dropped; a NULL outputs list is an empty list; NULL list children
// are skipped.  tweak_key: 64 bytes (the reference's format) or 33 bytes (compressed SEC1, decompressed in K1); a
// batch holds one format, so a change of key length cuts the batch.  Any other length can never match (the reference
// reads out of bounds there): dropped.
static void AccumulateInput(CudaspScanState &state, CudaspScanLocalState &local, const CudaspScanBindData &bind,
                            DataChunk &input) {
	if (input.ColumnCount() < 4) {
		throw InvalidInputException("cudasp_scan input must have 4 columns: txid, height, tweak_key, outputs");
	}
	auto &out_type = input.data[3].GetType();
	if (input.data[0].GetType().InternalType() != PhysicalType::VARCHAR || input.data[1].GetType().id() != LogicalTypeId::INTEGER ||
	    input.data[2].GetType().InternalType() != PhysicalType::VARCHAR || out_type.id() != LogicalTypeId::LIST ||
	    ListType::GetChildType(out_type).id() != LogicalTypeId::BIGINT) {
		throw InvalidInputException("cudasp_scan input columns must be (txid BLOB, height INTEGER, tweak_key BLOB, outputs BIGINT[])");
	}
	const idx_t count = input.size();
	UnifiedVectorFormat txid_f, height_f, tweak_f, list_f, child_f;
	input.data[0].ToUnifiedFormat(count, txid_f);
	input.data[1].ToUnifiedFormat(count, height_f);
	input.data[2].ToUnifiedFormat(count, tweak_f);
	input.data[3].ToUnifiedFormat(count, list_f);
	auto &child = ListVector::GetEntry(input.data[3]);
	child.ToUnifiedFormat(ListVector::GetListSize(input.data[3]), child_f);
	auto txids = UnifiedVectorFormat::GetData<string_t>(txid_f);
	auto heights = UnifiedVectorFormat::GetData<int32_t>(height_f);
	auto tweaks = UnifiedVectorFormat::GetData<string_t>(tweak_f);
	auto lists = UnifiedVectorFormat::GetData<list_entry_t>(list_f);
	auto values = UnifiedVectorFormat::GetData<int64_t>(child_f);

	// Fast path (what a Parquet or table scan delivers): no NULL in this chunk and the lists of consecutive rows adjacent in
	// the child vector.  Rows are then staged column by column for as many rows as fit the batch: one fixed-size copy per
	// key and per txid, one pass over the heights, ONE copy for the outputs of the whole run.  Dictionary-encoded Parquet
	// pages arrive as dictionary vectors (a selection vector over the dictionary, and a validity mask that exists even
	// when every entry is valid), so txid / height / tweak_key are read through their selection vectors and "no NULL" is
	// established by looking at the rows, not at the mask pointer.  Any row that breaks the pattern (other key length, a
	// gap between lists, outputs that overflow the slot) ends the run; the general row-by-row loop below takes over from
	// there for the rest of the chunk.
	auto all_rows_valid = [count](const UnifiedVectorFormat &f) {
		if (f.validity.AllValid()) {
			return true;
		}
		for (idx_t i = 0; i < count; i++) {
			if (!f.validity.RowIsValid(f.sel->get_index(i))) {
				return false;
			}
		}
		return true;
	};
	idx_t r = 0;
	const bool flat = !list_f.sel->IsSet() && !child_f.sel->IsSet() && list_f.validity.AllValid() && child_f.validity.AllValid() &&
	                  all_rows_valid(txid_f) && all_rows_valid(height_f) && all_rows_valid(tweak_f);
	const sel_t *tsel = txid_f.sel->IsSet() ? txid_f.sel->data() : nullptr;
	const sel_t *hsel = height_f.sel->IsSet() ? height_f.sel->data() : nullptr;
	const sel_t *wsel = tweak_f.sel->IsSet() ? tweak_f.sel->data() : nullptr;
#define CUDASP_AT(sel, i) ((sel) ? (idx_t)(sel)[i] : (idx_t)(i))
	while (flat && r < count) {
		const idx_t key_size = tweaks[CUDASP_AT(wsel, r)].GetSize();
		if (key_size != POINT_SIZE && key_size != SEC1_SIZE) {
			break;
		}
		if (local.have_slot && local.rows && local.key_bytes != key_size) {
			SubmitSlot(state, local);
		}
		EnsureSlot(state, local);
		if (local.rows == 0) {
			local.key_bytes = local.meta->key_bytes = (uint32_t)key_size;
		}
		const uint32_t row0 = local.rows;
		idx_t n = MinValue<idx_t>(count - r, MinValue<uint32_t>(local.batch_limit, local.slot.cap_rows) - row0);
		// lists: adjacent, and fitting the slot's output capacity
		const idx_t first_off = lists[r].offset;
		uint64_t acc = 0;
		uint32_t *offs = local.slot.out_offsets + row0 + 1;
		const uint64_t room = local.slot.cap_outputs - local.n_outputs;
		idx_t m = 0;
		for (; m < n; m++) {
			const list_entry_t &le = lists[r + m];
			if (le.offset != first_off + acc || acc + le.length > room) {
				break;
			}
			acc += le.length;
			offs[m] = (uint32_t)(local.n_outputs + acc);
		}
		n = m;
		// keys: one length for the run
		uint8_t *kdst = local.slot.tweak_keys + (size_t)row0 * key_size;
		if (key_size == POINT_SIZE) {
			for (m = 0; m < n; m++) {
				const string_t &tw = tweaks[CUDASP_AT(wsel, r + m)];
				if (tw.GetSize() != POINT_SIZE) {
					break;
				}
				CopyStreaming<POINT_SIZE>(kdst + m * POINT_SIZE, tw.GetData());
			}
		} else {
			for (m = 0; m < n; m++) {
				const string_t &tw = tweaks[CUDASP_AT(wsel, r + m)];
				if (tw.GetSize() != SEC1_SIZE) {
					break;
				}
				memcpy(kdst + m * SEC1_SIZE, tw.GetData(), SEC1_SIZE);
			}
		}
		n = m;
		if (n == 0) {
			break;  // the row at r does not fit this pattern (e.g. its outputs overflow an empty slot): general path
		}
		const uint64_t n_out = (uint64_t)offs[n - 1] - local.n_outputs;
		if (n_out) {
			memcpy(local.slot.outputs + local.n_outputs, values + first_off, n_out * sizeof(int64_t));
		}
		auto &meta = *local.meta;
		int32_t *hdst = meta.heights.Grow(n);
		if (hsel) {
			for (m = 0; m < n; m++) {
				hdst[m] = heights[hsel[r + m]];
			}
		} else {
			memcpy(hdst, heights + r, n * sizeof(int32_t));
		}
		if (meta.txid_fixed) {
			char *tdst = meta.txid_bytes.Grow(n * TXID_SIZE);
			for (m = 0; m < n; m++) {
				const string_t &tx = txids[CUDASP_AT(tsel, r + m)];
				if (tx.GetSize() != TXID_SIZE) {
					break;
				}
				CopyStreaming<TXID_SIZE>(tdst + m * TXID_SIZE, tx.GetData());
			}
			if (m < n) {  // an odd-sized txid: switch this batch to the explicit index and append the rest one by one
				meta.txid_bytes.Truncate((size_t)(row0 + m) * TXID_SIZE);
				meta.rows = (uint32_t)(row0 + m);
				for (; m < n; m++) {
					meta.AddTxid(txids[CUDASP_AT(tsel, r + m)], TXID_SIZE);
					meta.rows++;
				}
			}
		} else {
			for (m = 0; m < n; m++) {
				meta.AddTxid(txids[CUDASP_AT(tsel, r + m)], TXID_SIZE);
			}
		}
		meta.rows = (uint32_t)(row0 + n);
		local.n_outputs += n_out;
		local.rows = (uint32_t)(row0 + n);
		local.rows_fast += n;
		r += n;
		if (local.rows >= local.batch_limit || local.rows >= local.slot.cap_rows) {
			SubmitSlot(state, local);
		}
	}
#undef CUDASP_AT

	for (; r < count; r++) {
		const idx_t ti = txid_f.sel->get_index(r), hi = height_f.sel->get_index(r), wi = tweak_f.sel->get_index(r);
		if (!txid_f.validity.RowIsValid(ti) || !height_f.validity.RowIsValid(hi) || !tweak_f.validity.RowIsValid(wi)) {
			continue;
		}
		const string_t &tw = tweaks[wi];
		const idx_t key_size = tw.GetSize();
		if (key_size != POINT_SIZE && key_size != SEC1_SIZE) {
			continue;
		}
		if (local.have_slot && local.rows && local.key_bytes != key_size) {
			SubmitSlot(state, local);
		}
		const idx_t li = list_f.sel->get_index(r);
		const bool has_list = list_f.validity.RowIsValid(li);
		const idx_t list_len = has_list ? lists[li].length : 0;
		EnsureSlot(state, local);
		// cut the batch early if this row's outputs would overflow the slot's output capacity
		if (local.rows && local.n_outputs + list_len > local.slot.cap_outputs) {
			SubmitSlot(state, local);
			EnsureSlot(state, local);
		}
		if (list_len > local.slot.cap_outputs) {
			throw InvalidInputException("cudasp_scan: a row has %llu outputs, more than a batch can hold (%llu)",
			                            (uint64_t)list_len, (uint64_t)local.slot.cap_outputs);
		}
		const uint32_t row = local.rows;
		if (row == 0) {
			local.key_bytes = local.meta->key_bytes = (uint32_t)key_size;
		}
		memcpy(local.slot.tweak_keys + (size_t)row * key_size, tw.GetData(), key_size);
		uint64_t n_out = local.n_outputs;
		if (has_list) {
			const idx_t off = lists[li].offset;
			for (idx_t j = 0; j < list_len; j++) {
				const idx_t ci = child_f.sel->get_index(off + j);
				if (child_f.validity.RowIsValid(ci)) {
					local.slot.outputs[n_out++] = values[ci];
				}
			}
		}
		local.n_outputs = n_out;
		local.slot.out_offsets[row + 1] = (uint32_t)n_out;
		auto &meta = *local.meta;
		meta.AddTxid(txids[ti], TXID_SIZE);
		*meta.heights.Grow(1) = heights[hi];
		meta.rows = row + 1;
		local.rows = row + 1;
		if (local.rows >= local.batch_limit || local.rows >= local.slot.cap_rows) {
			SubmitSlot(state, local);
		}
	}
}

// Copy up to STANDARD_VECTOR_SIZE collected result rows into `output` (under the output lock: other workers may be
// appending).  Returns true when more rows are waiting.
static bool EmitRows(CudaspScanState &state, DataChunk &output) {
	std::lock_guard<std::mutex> guard(state.output_lock);
	const idx_t remaining = state.output_txids.size() - state.output_position;
	const idx_t n = MinValue<idx_t>(STANDARD_VECTOR_SIZE, remaining);
	if (n > 0) {
		auto txid_out = FlatVector::GetData<string_t>(output.data[0]);
		auto height_out = FlatVector::GetData<int32_t>(output.data[1]);
		auto tweak_out = FlatVector::GetData<string_t>(output.data[2]);
		for (idx_t i = 0; i < n; i++) {
			const idx_t src = state.output_position + i;
			txid_out[i] = StringVector::AddStringOrBlob(output.data[0], state.output_txids[src]);
			height_out[i] = state.output_heights[src];
			tweak_out[i] = StringVector::AddStringOrBlob(output.data[2], state.output_tweak_keys[src]);
		}
		if (state.details) {
			auto k_out = FlatVector::GetData<uint32_t>(output.data[3]);
			auto oi_out = FlatVector::GetData<uint32_t>(output.data[4]);
			auto li_out = FlatVector::GetData<int32_t>(output.data[5]);
			auto t_out = FlatVector::GetData<string_t>(output.data[6]);
			for (idx_t i = 0; i < n; i++) {
				const cudasp_match_detail &d = state.output_details[state.output_position + i];
				k_out[i] = d.k;
				oi_out[i] = d.output_index;
				li_out[i] = d.label_index;
				t_out[i] = StringVector::AddStringOrBlob(output.data[6], const_char_ptr_cast(d.t_k), 32);
			}
		}
		output.SetCardinality(n);
		state.output_position += n;
	}
	return state.output_position < state.output_txids.size();
}

// By default never emits during Execute (same protocol as the reference, :534-552); finished batches are collected
// opportunistically so that slots recycle while the scan is still feeding rows.
static OperatorResultType CudaspScanFunction(ExecutionContext &context, TableFunctionInput &data_p, DataChunk &input,
                                             DataChunk &output) {
	auto &bind = data_p.bind_data->Cast<CudaspScanBindData>();
	auto &state = data_p.global_state->Cast<CudaspScanState>();
	auto &local = data_p.local_state->Cast<CudaspScanLocalState>();
	if (input.size() > 0 && !local.input_consumed) {
		const double t0 = TraceOn() ? NowSec() : 0;
		if (TraceOn() && local.rows_seen == 0) {
			local.first_input = t0;
		}
		AccumulateInput(state, local, bind, input);
		// Finished batches are collected when a worker needs a slot (EnsureSlot) and at the end (Final).  Polling after
		// every 2 048-row chunk from every worker only contends for the query's lock; with emit_early it is what
		// brings the matches out early, so it stays there (every fourth chunk of a worker).
		if (bind.emit_early && (++local.chunks_since_poll & 3) == 0) {
			Drain(state, false);
		}
		if (TraceOn()) {
			local.accumulate_s += NowSec() - t0;
			local.rows_seen += input.size();
		}
	}
	if (bind.emit_early) {
		// extension: hand finished matches downstream now instead of in Final.  When a full
		// vector went out and more is waiting, DuckDB calls again with the SAME input: it must not be staged twice.
		if (EmitRows(state, output) && output.size() == STANDARD_VECTOR_SIZE) {
			local.input_consumed = true;
			return OperatorResultType::HAVE_MORE_OUTPUT;
		}
	}
	local.input_consumed = false;
	return OperatorResultType::NEED_MORE_INPUT;
}

// ------------------------------------------------------------------------------------------ finalize
// Same protocol as the reference's CudaspScanFinalFunction (:554-624): every worker flushes its partial batch; the
// LAST worker to finalize claims the output role, waits for everything in flight and streams the global result out
// STANDARD_VECTOR_SIZE rows at a time.  (One emitting thread keeps DuckDB's batch-index bookkeeping happy, :580-589.)
static OperatorFinalizeResultType CudaspScanFinalFunction(ExecutionContext &context, TableFunctionInput &data_p,
                                                          DataChunk &output) {
	auto &state = data_p.global_state->Cast<CudaspScanState>();
	auto &local = data_p.local_state->Cast<CudaspScanLocalState>();

	if (!local.finalized) {
		if (TraceOn()) {
			fprintf(stderr, "[cudasp] worker: %llu rows (%llu column-wise), accumulate+drain %.3f s (of which waiting for a slot %.3f s), span %.3f s\n",
			        (unsigned long long)local.rows_seen, (unsigned long long)local.rows_fast, local.accumulate_s, local.slot_wait_s,
			        local.rows_seen ? NowSec() - local.first_input : 0.0);
		}
		SubmitSlot(state, local);
		Drain(state, false);
		std::lock_guard<std::mutex> guard(state.finalize_lock);
		state.currently_adding--;
		local.finalized = true;
	}
	if (!local.is_output_thread) {
		if (state.currently_adding != 0) {
			return OperatorFinalizeResultType::FINISHED;
		}
		bool expected = false;
		if (!state.output_thread_claimed.compare_exchange_strong(expected, true)) {
			return OperatorFinalizeResultType::FINISHED;
		}
		local.is_output_thread = true;
		const double t0 = NowSec();
		Drain(state, true);  // barrier: every submitted batch has been collected after this
		if (TraceOn()) {
			fprintf(stderr, "[cudasp] final drain: %.3f s, %llu matches\n", NowSec() - t0, (unsigned long long)state.output_txids.size());
		}
	}

	EmitRows(state, output);
	return state.output_position < state.output_txids.size() ? OperatorFinalizeResultType::HAVE_MORE_OUTPUT
	                                                         : OperatorFinalizeResultType::FINISHED;
}

// ------------------------------------------------------------------------------------------ cudasp_scan_parquet
// Sibling of cudasp_scan for tables that live in Parquet files (no counterpart in the reference, whose users go
// through read_parquet + AccumulateInput, README.md:100-110):
//     SELECT * FROM cudasp_scan_parquet('tweaks.parquet' | ['a.parquet', ...], scan_private_key, spend_public_key, label_keys
//                                       [, batch_size := N] [, threads := T])
// Same key arguments, checks, error texts and result columns as cudasp_scan.  The pages are decoded by the C-ABI library itself
// (include/cudasp_parquet.h) straight into pinned staging, one row group per host thread in flight, without becoming DuckDB vectors.
// Files outside that reader's subset (codecs other than none / Snappy, other column shapes) are refused with a message that names
// the feature and points at cudasp_scan((SELECT ... FROM read_parquet(...)), ...), which reads anything DuckDB can read.
struct CudaspParquetBindData : public CudaspScanBindData {
	vector<string> paths;
	int64_t threads = 0;
};

struct CudaspParquetState : public GlobalTableFunctionState {
	~CudaspParquetState() override {
		if (result) {
			cudasp_parquet_result_free(result);
		}
		if (query) {
			cudasp_query_destroy(query);
		}
	}
	idx_t MaxThreads() const override {
		return 1;  // the library runs its own decoder threads
	}
	cudasp_query *query = nullptr;
	cudasp_parquet_result *result = nullptr;
	bool scanned = false;
	idx_t position = 0;
};

static unique_ptr<FunctionData> CudaspParquetBind(ClientContext &context, TableFunctionBindInput &input,
                                                  vector<LogicalType> &return_types, vector<string> &names) {
	if (input.inputs.size() != 4) {
		throw InvalidInputException("cudasp_scan_parquet requires 4 arguments: path VARCHAR or LIST[VARCHAR], scan_private_key "
		                            "BLOB, spend_public_key BLOB, and label_keys LIST[BLOB]");
	}
	auto result = make_uniq<CudaspParquetBindData>();
	auto &p = input.inputs[0];
	if (p.IsNull()) {
		throw InvalidInputException("First argument must be a VARCHAR path or a LIST[VARCHAR] of paths, not NULL");
	}
	if (p.type().id() == LogicalTypeId::VARCHAR) {
		result->paths.push_back(StringValue::Get(p));
	} else if (p.type().id() == LogicalTypeId::LIST) {
		for (auto &child : ListValue::GetChildren(p)) {
			if (child.type().id() != LogicalTypeId::VARCHAR || child.IsNull()) {
				throw InvalidInputException("All elements of the path list must be VARCHARs");
			}
			result->paths.push_back(StringValue::Get(child));
		}
	} else {
		throw InvalidInputException("First argument must be a VARCHAR path or a LIST[VARCHAR] of paths");
	}
	// glob patterns as in read_parquet('dir/*.parquet'): expanded here, in sorted order; the library maps LOCAL files, so a path
	// another file system would serve (s3://, http://) is refused by the scan with the read_parquet form to use instead
	{
		auto &fs = FileSystem::GetFileSystem(context);
		vector<string> expanded;
		for (auto &path : result->paths) {
			if (!FileSystem::HasGlob(path)) {
				expanded.push_back(path);
				continue;
			}
			auto found = fs.GlobFiles(path, context, FileGlobOptions::DISALLOW_EMPTY);
			std::sort(found.begin(), found.end());
			for (auto &info : found) {
				expanded.push_back(info.path);
			}
		}
		result->paths = std::move(expanded);
	}
	BindKeys(input, *result);
	auto th = input.named_parameters.find("threads");
	if (th != input.named_parameters.end() && !th->second.IsNull()) {
		result->threads = th->second.GetValue<int64_t>();
		if (result->threads < 0 || result->threads > 1024) {
			throw InvalidInputException("threads must be between 0 and 1024, got %lld", (long long)result->threads);
		}
	}
	return_types = {LogicalType::BLOB, LogicalType::INTEGER, LogicalType::BLOB};
	names = {"txid", "height", "tweak_key"};
	return std::move(result);
}

static unique_ptr<GlobalTableFunctionState> CudaspParquetInit(ClientContext &context, TableFunctionInitInput &input) {
	return make_uniq<CudaspParquetState>();
}

static void CudaspParquetFunction(ClientContext &context, TableFunctionInput &data_p, DataChunk &output) {
	auto &bind = data_p.bind_data->Cast<CudaspParquetBindData>();
	auto &state = data_p.global_state->Cast<CudaspParquetState>();
	if (!state.scanned) {
		state.scanned = true;
		int n_dev = cudasp_device_count();
		if (n_dev <= 0) {
			throw IOException("cudasp_scan_parquet: no CUDA device available (this extension has no CPU path)");
		}
		idx_t threads = bind.threads > 0 ? (idx_t)bind.threads
		                                 : MaxValue<idx_t>(1, NumericCast<idx_t>(TaskScheduler::GetScheduler(context).NumberOfThreads()));
		// every decoder thread holds one slot while it fills it (pieces of at most batch_size rows, cut across row groups); four
		// more per device keep copies and batches in flight behind them
		uint32_t slots = (uint32_t)MinValue<idx_t>(64, (threads + n_dev - 1) / n_dev + 4);
		const double t_create = NowSec();
		int rc = cudasp_query_create_ex(const_data_ptr_cast(bind.scan_private_key.data()), const_data_ptr_cast(bind.spend_public_key.data()),
		                                bind.n_labels ? const_data_ptr_cast(bind.label_keys.data()) : nullptr, bind.n_labels, nullptr, 0,
		                                (uint32_t)bind.batch_size, 8ull * bind.batch_size, slots, &state.query);
		if (rc != CUDASP_OK) {
			const char *msg = cudasp_last_error(nullptr);
			if (rc == CUDASP_ERR_KEY || rc == CUDASP_ERR_ARG) {
				throw InvalidInputException("cudasp_scan_parquet: %s", msg ? msg : "invalid keys");
			}
			throw IOException("cudasp_scan_parquet: GPU initialisation failed: %s", msg ? msg : "unknown error");
		}
		vector<const char *> cpaths;
		for (auto &p : bind.paths) {
			cpaths.push_back(p.c_str());
		}
		const double t_scan = NowSec();
		rc = cudasp_scan_parquet(state.query, cpaths.data(), (int)cpaths.size(), (int)threads, &state.result);
		if (getenv("CUDASP_TRACE")) {
			fprintf(stderr, "[cudasp] cudasp_scan_parquet: query_create %.3f s (%u slots/device, %d devices, %llu threads), scan %.3f s\n",
			        t_scan - t_create, slots, n_dev, (unsigned long long)threads, NowSec() - t_scan);
		}
		if (rc != CUDASP_OK) {
			const char *msg = cudasp_last_error(state.query);
			if (rc == CUDASP_ERR_UNSUPPORTED) {
				throw InvalidInputException("cudasp_scan_parquet: %s -- use cudasp_scan((SELECT txid, height, tweak_key, outputs FROM "
				                            "read_parquet(...)), ...) for this file",
				                            msg ? msg : "unsupported Parquet feature");
			}
			throw IOException("cudasp_scan_parquet: %s", msg ? msg : "scan failed");
		}
	}
	const cudasp_parquet_result &r = *state.result;
	const idx_t n = MinValue<idx_t>(STANDARD_VECTOR_SIZE, (idx_t)r.n_matches - state.position);
	auto txid_out = FlatVector::GetData<string_t>(output.data[0]);
	auto height_out = FlatVector::GetData<int32_t>(output.data[1]);
	auto tweak_out = FlatVector::GetData<string_t>(output.data[2]);
	for (idx_t i = 0; i < n; i++) {
		const idx_t k = state.position + i;
		txid_out[i] = StringVector::AddStringOrBlob(output.data[0], const_char_ptr_cast(r.txids + r.txid_offsets[k]),
		                                            (idx_t)(r.txid_offsets[k + 1] - r.txid_offsets[k]));
		height_out[i] = r.heights[k];
		tweak_out[i] = StringVector::AddStringOrBlob(output.data[2], const_char_ptr_cast(r.tweak_keys + r.key_offsets[k]),
		                                             (idx_t)(r.key_offsets[k + 1] - r.key_offsets[k]));
	}
	output.SetCardinality(n);
	state.position += n;
}

// ------------------------------------------------------------------------------------------ registration
static void LoadInternal(ExtensionLoader &loader) {
	TableFunction func("cudasp_scan",
	                   {LogicalType::TABLE, LogicalType::BLOB, LogicalType::BLOB, LogicalType::LIST(LogicalType::BLOB)}, nullptr,
	                   CudaspScanBind, CudaspScanInit, CudaspScanLocalInit);
	func.in_out_function = CudaspScanFunction;
	func.in_out_function_final = CudaspScanFinalFunction;
	func.named_parameters["batch_size"] = LogicalType::INTEGER;
	func.named_parameters["details"] = LogicalType::BOOLEAN;
	func.named_parameters["emit_early"] = LogicalType::BOOLEAN;
	loader.RegisterFunction(func);

	for (auto &path_type : {LogicalType(LogicalType::VARCHAR), LogicalType::LIST(LogicalType::VARCHAR)}) {
		TableFunction pfunc("cudasp_scan_parquet", {path_type, LogicalType::BLOB, LogicalType::BLOB, LogicalType::LIST(LogicalType::BLOB)},
		                    CudaspParquetFunction, CudaspParquetBind, CudaspParquetInit);
		pfunc.named_parameters["batch_size"] = LogicalType::INTEGER;
		pfunc.named_parameters["threads"] = LogicalType::INTEGER;
		loader.RegisterFunction(pfunc);
	}
}

void CudaspExtension::Load(ExtensionLoader &loader) {
	LoadInternal(loader);
}
std::string CudaspExtension::Name() {
	return "cudasp";
}
std::string CudaspExtension::Version() const {
#ifdef EXT_VERSION_CUDASP
	return EXT_VERSION_CUDASP;
#else
	return cudasp_version();
#endif
}

}  // namespace duckdb

extern "C" {
DUCKDB_CPP_EXTENSION_ENTRY(cudasp, loader) {
	duckdb::LoadInternal(loader);
}
}
