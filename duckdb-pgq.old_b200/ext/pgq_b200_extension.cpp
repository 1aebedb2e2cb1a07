// pgq_b200.duckdb_extension — DuckDB host shim of the B200 graph hot path.
//
// Registers the DuckPGQ scalar functions (same SQL names, argument vectors and result vectors;
// SURVEY.md §8(a) a2-a7) through DuckDB's scalar-function API and forwards every chunk to the
// C ABI in include/pgq_b200.h.  Nothing here computes: no CPU fallback exists, a missing GPU
// or library surfaces as a duckdb::Exception with the C ABI's message.
//
// Contracts followed (all in /root/reference):
//   plugin ABI  <basename>_init / <basename>_version   src/execution/operator/helper/physical_load.cpp:36-65
//   registration via Catalog::CreateFunction            extension/fts/fts-extension.cpp:51-74
//   scalar_function_t / bind_scalar_function_t          src/include/duckdb/function/scalar_function.hpp:24-27
//   one call = one DataChunk of <= 1024 rows            src/execution/expression_executor/execute_function.cpp:19-40
//   Orrify / VectorData                                 src/include/duckdb/common/types/vector.hpp:22-27,116-119
//   LIST result fill (Reserve / GetEntry / SetListSize) src/function/scalar/list/range.cpp:118-149
//   CONSTANT result for all-constant inputs             src/function/scalar/list/range.cpp:105-152
//   NULL in / out                                       src/include/duckdb/common/types/validity_mask.hpp:151-156,190-196
//   has_side_effects = true (shared state)              src/function/scalar/sequence/nextval.cpp:137-138
//   concurrent entry from PRAGMA threads workers        src/parallel/pipeline.cpp:79-123
//
// Concurrency.  The engine enters a function from up to `PRAGMA threads` workers at once,
// one <= 1024-row chunk each.  A CSR slot has a reader/writer lock (builds and delete exclusive, queries shared)
// and a cross-call batcher: query calls of different workers queue their rows; the first caller becomes the
// leader, hands ALL queued rows of its kind to one C-ABI call (so the GPU sees whole 512-lane batches and
// several of them in flight instead of one 1024-row call at a time) and wakes the others with their slices.
// Blocking inside a scalar function is what the reference's own stateful functions do (nextval takes
// seq->lock: src/function/scalar/sequence/nextval.cpp:31,43).
#define DUCKDB_BUILD_LOADABLE_EXTENSION
#include "duckdb.hpp"
#include "duckdb/catalog/catalog.hpp"
#include "duckdb/common/exception.hpp"
#include "duckdb/function/scalar_function.hpp"
#include "duckdb/main/client_context.hpp"
#include "duckdb/parser/parsed_data/create_scalar_function_info.hpp"
#include "duckdb/planner/expression/bound_function_expression.hpp"

#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <vector>

#include "../../include/pgq_b200.h"
#include "pgq_rewrite.hpp"

namespace duckdb {

// ---- small reader/writer lock (C++11: no std::shared_mutex) --------------------------------------------
struct PgqRwLock {
	std::mutex m;
	std::condition_variable cv;
	int readers = 0;
	bool writer = false;
	void LockShared() {
		std::unique_lock<std::mutex> l(m);
		cv.wait(l, [&] { return !writer; });
		readers++;
	}
	void UnlockShared() {
		std::lock_guard<std::mutex> l(m);
		if (--readers == 0) cv.notify_all();
	}
	void Lock() {
		std::unique_lock<std::mutex> l(m);
		cv.wait(l, [&] { return !writer && readers == 0; });
		writer = true;
	}
	void Unlock() {
		std::lock_guard<std::mutex> l(m);
		writer = false;
		cv.notify_all();
	}
};
struct PgqSharedGuard {
	PgqRwLock &l;
	explicit PgqSharedGuard(PgqRwLock &l) : l(l) {
		l.LockShared();
	}
	~PgqSharedGuard() {
		l.UnlockShared();
	}
};
struct PgqExclusiveGuard {
	PgqRwLock &l;
	explicit PgqExclusiveGuard(PgqRwLock &l) : l(l) {
		l.Lock();
	}
	~PgqExclusiveGuard() {
		l.Unlock();
	}
};

// ---- one query call waiting in a slot's batcher -----------------------------------------------------------
enum PgqKind { PGQ_K_LENGTH = 0, PGQ_K_REACH, PGQ_K_BOUNDED, PGQ_K_PATH0, PGQ_K_PATH1, PGQ_K_CHEAP_I64, PGQ_K_CHEAP_F64 };

struct PgqPending {
	int kind = 0;
	int64_t lower = 0, upper = -1;
	idx_t n = 0;
	const int64_t *src = nullptr, *dst = nullptr;
	const uint8_t *sv = nullptr, *dv = nullptr;
	// results (caller-owned arrays of n entries)
	void *out8 = nullptr;    // lengths (int64) / cheapest path lengths (int64 or double)
	uint8_t *outb = nullptr; // reachability
	uint8_t *ov = nullptr;
	pgq_list_entry_t *entries = nullptr;            // paths: offsets relative to this call's slice of the values
	std::shared_ptr<std::vector<int64_t>> vals_all; // paths: the leader's combined value buffer
	uint64_t vals_begin = 0, vals_count = 0;        // this call's slice of it
	bool done = false;
	int rc = PGQ_OK;
	std::string err;
	bool SameGroup(const PgqPending &o) const {
		return kind == o.kind && lower == o.lower && upper == o.upper;
	}
};

// ---- process-wide GPU context + (connection, csr_id) -> device csr id -----------------------------
struct PgqCsrSlot {
	int32_t dev_id = -1;
	std::weak_ptr<ClientContext> owner; // the connection the slot belongs to (R1: per-connection map)
	bool vertex_started = false; // CsrInitializeVertex done (first create_csr_vertex row seen)
	bool edge_started = false;   // CsrInitializeEdge done (first create_csr_edge row seen)
	bool finalized = false;
	bool dropped = false; // delete_csr ran while a caller still held the slot
	int64_t num_edges = 0, rows_added = 0;
	// 7-argument create_csr_edge: weights of the rows seen so far (raw 8-byte values, input-row order)
	int wtype = 0;
	std::vector<uint64_t> weights;
	PgqRwLock rw; // builds / finalize / delete exclusive, queries shared
	// cross-call batcher
	std::mutex bm;
	std::condition_variable bcv;
	std::vector<PgqPending *> queue;
	bool leader_active = false;
	uint64_t calls = 0, gpu_calls = 0; // statistics: query calls entered / C-ABI calls issued for them
};

struct PgqState {
	std::mutex mu; // the slot map and the lazily created GPU context; never held across a GPU call
	pgq_ctx_t *ctx = nullptr;
	int32_t next_dev_id = 1;
	std::map<std::pair<ClientContext *, int32_t>, std::shared_ptr<PgqCsrSlot>> slots;

	[[noreturn]] static void Fail(const char *what, int rc, const char *msg = nullptr) {
		if (!msg) msg = pgq_last_error();
		string text = string("pgq_b200: ") + what + ": " + (msg && *msg ? msg : "error") + " (code " +
		              std::to_string(rc) + ")";
		if (rc == PGQ_ERR_OOM) throw OutOfMemoryException(text);
		if (rc == PGQ_ERR_INVALID || rc == PGQ_ERR_MISMATCH) throw InvalidInputException(text);
		throw InternalException(text);
	}
	static void Check(const char *what, int rc) {
		if (rc != PGQ_OK) Fail(what, rc);
	}
	// caller holds mu.  PGQ_B200_DEVICES=0,1,2,... : one multi-GPU context (pgq_init_devices): queries deal their
	// batches over all listed GPUs; PGQ_B200_DEVICE=n : one GPU (default 0)
	pgq_ctx_t *CtxLocked() {
		if (!ctx) {
			const char *devs = std::getenv("PGQ_B200_DEVICES");
			if (devs && *devs) {
				std::vector<int> list;
				const char *p = devs;
				while (*p) {
					char *end = nullptr;
					long v = std::strtol(p, &end, 10);
					if (end == p) break;
					list.push_back((int)v);
					p = (*end == ',') ? end + 1 : end;
				}
				Check("pgq_init_devices", pgq_init_devices(list.data(), (int)list.size(), &ctx));
			} else {
				const char *dev = std::getenv("PGQ_B200_DEVICE");
				Check("pgq_init", pgq_init(dev ? std::atoi(dev) : 0, &ctx));
			}
		}
		return ctx;
	}
	// caller holds mu: slots of connections that no longer exist are dropped (their device CSRs freed), so a
	// ClientContext allocated at a recycled address never inherits a dead connection's graph
	void PurgeLocked() {
		for (auto it = slots.begin(); it != slots.end();) {
			if (it->second->owner.expired()) {
				if (it->second->vertex_started && ctx) pgq_csr_delete(ctx, it->second->dev_id);
				it->second->dropped = true;
				it = slots.erase(it);
			} else {
				++it;
			}
		}
	}
	std::shared_ptr<PgqCsrSlot> Slot(ClientContext *cc, int32_t csr_id, bool create) {
		std::lock_guard<std::mutex> lock(mu);
		CtxLocked();
		PurgeLocked();
		auto key = std::make_pair(cc, csr_id);
		auto it = slots.find(key);
		if (it == slots.end()) {
			if (!create) return nullptr;
			auto s = std::make_shared<PgqCsrSlot>();
			s->dev_id = next_dev_id++;
			s->owner = cc->shared_from_this();
			it = slots.insert(std::make_pair(key, s)).first;
		}
		return it->second;
	}
};

static PgqState &State() {
	static PgqState st;
	return st;
}

// caller holds the slot's exclusive lock: K1-K3 run now; weights (if the rows carried any) go into the CSR slots
static void FinishSlot(pgq_ctx_t *pctx, PgqCsrSlot &s) {
	if (!s.edge_started) { // vertices but no edge row at all: an empty edge set still finalizes
		PgqState::Check("create_csr_edge", pgq_csr_begin_edges(pctx, s.dev_id, 0));
		s.edge_started = true;
		s.num_edges = 0;
	}
	PgqState::Check("csr finalize", pgq_csr_finalize(pctx, s.dev_id));
	s.finalized = true;
	if (s.wtype) {
		PgqState::Check("csr weights", pgq_csr_set_weights(pctx, s.dev_id, s.wtype, s.weights.data(), 0,
		                                                   (int64_t)s.weights.size()));
		s.weights.clear();
		s.weights.shrink_to_fit();
	}
}

// caller holds the slot's exclusive lock: forget a half-built CSR so the next create_csr_* starts clean
static void ResetSlot(pgq_ctx_t *pctx, PgqCsrSlot &s) {
	if (s.vertex_started) pgq_csr_delete(pctx, s.dev_id);
	s.vertex_started = s.edge_started = s.finalized = false;
	s.rows_added = s.num_edges = 0;
	s.wtype = 0;
	s.weights.clear();
}

// the slot of a query: must exist; finalized lazily when the edge rows stopped short of num_edges
static std::shared_ptr<PgqCsrSlot> ReadySlot(ClientContext *cc, int32_t csr_id) {
	auto &st = State();
	auto slot = st.Slot(cc, csr_id, false);
	if (!slot || !slot->vertex_started) {
		throw InvalidInputException("pgq_b200: csr " + std::to_string(csr_id) +
		                            " does not exist (run create_csr_vertex / create_csr_edge first)");
	}
	if (!slot->finalized) {
		PgqExclusiveGuard g(slot->rw);
		if (!slot->finalized && slot->vertex_started) FinishSlot(st.ctx, *slot);
	}
	return slot;
}

// bind data: the connection the CSR map is keyed by (cf. NextvalBindData capturing the context,
// src/function/scalar/sequence/nextval.cpp:15-27)
struct PgqBindData : public FunctionData {
	explicit PgqBindData(ClientContext &context) : context(context) {
	}
	ClientContext &context;
	unique_ptr<FunctionData> Copy() override {
		return make_unique<PgqBindData>(context);
	}
};

static unique_ptr<FunctionData> PgqBind(ClientContext &context, ScalarFunction &, vector<unique_ptr<Expression>> &) {
	return make_unique<PgqBindData>(context);
}

static ClientContext *ContextOf(ExpressionState &state) {
	auto &expr = (BoundFunctionExpression &)state.expr;
	return &((PgqBindData &)*expr.bind_info).context;
}

// ---- argument helpers --------------------------------------------------------------------------------
// constant-per-call arguments (csr_id, v_size, num_edges) are read from row 0 like the reference does
static int64_t FirstBigint(DataChunk &args, idx_t col) {
	VectorData vd;
	args.data[col].Orrify(args.size(), vd);
	auto idx = vd.sel->get_index(0);
	if (!vd.validity.RowIsValid(idx)) throw InvalidInputException("pgq_b200: NULL csr_id / size argument");
	return ((int64_t *)vd.data)[idx];
}
static int32_t FirstInt(DataChunk &args, idx_t col) {
	VectorData vd;
	args.data[col].Orrify(args.size(), vd);
	auto idx = vd.sel->get_index(0);
	if (!vd.validity.RowIsValid(idx)) throw InvalidInputException("pgq_b200: NULL csr_id argument");
	return ((int32_t *)vd.data)[idx];
}

// every input vector is a constant: the result is one too, computed from row 0 (range.cpp:105-152 pattern).
// Only the read-only query functions use this; the create_csr_* functions count every row they are given.
static bool AllConstant(DataChunk &args) {
	for (idx_t c = 0; c < args.ColumnCount(); c++)
		if (args.data[c].GetVectorType() != VectorType::CONSTANT_VECTOR) return false;
	return true;
}

// dense copy of a BIGINT column + one validity byte per row
struct DenseCol {
	std::vector<int64_t> val;
	std::vector<uint8_t> valid;
	bool all_valid = true;
	void Load(Vector &v, idx_t count) {
		VectorData vd;
		v.Orrify(count, vd);
		val.resize(count);
		valid.resize(count);
		auto data = (int64_t *)vd.data;
		for (idx_t i = 0; i < count; i++) {
			auto idx = vd.sel->get_index(i);
			bool ok = vd.validity.RowIsValid(idx);
			valid[i] = ok;
			val[i] = ok ? data[idx] : 0;
			all_valid &= ok;
		}
	}
};

// ---- create_csr_vertex(INTEGER csr_id, BIGINT v_size, BIGINT vertex, BIGINT cnt) -> BIGINT (R2) ----
static void CreateCsrVertexFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	const idx_t count = args.size();
	result.SetVectorType(VectorType::FLAT_VECTOR);
	if (count == 0) return;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	const int64_t v_size = FirstBigint(args, 1);
	DenseCol vid, cnt;
	vid.Load(args.data[2], count);
	cnt.Load(args.data[3], count);
	auto out = FlatVector::GetData<int64_t>(result);
	auto &validity = FlatVector::Validity(result);
	// rows with a NULL vertex id or count carry no degree: NULL out, not forwarded
	std::vector<int64_t> fv, fc;
	fv.reserve(count);
	fc.reserve(count);
	for (idx_t i = 0; i < count; i++) {
		if (!vid.valid[i] || !cnt.valid[i]) {
			validity.SetInvalid(i);
			out[i] = 0;
			continue;
		}
		fv.push_back(vid.val[i]);
		fc.push_back(cnt.val[i]);
		out[i] = cnt.val[i];
	}
	auto slot = st.Slot(ContextOf(state), csr_id, true);
	PgqExclusiveGuard guard(slot->rw);
	auto *pctx = st.ctx;
	// First row of a (new) CSR: CsrInitializeVertex.  Vertex rows always precede the edge rows of a build, so a
	// vertex row on a slot whose edge phase has started (finished or aborted part-way) begins a NEW build.
	if (!slot->vertex_started || slot->finalized || slot->edge_started) {
		ResetSlot(pctx, *slot);
		PgqState::Check("create_csr_vertex", pgq_csr_create(pctx, slot->dev_id, v_size));
		slot->vertex_started = true;
	}
	int64_t sum = 0;
	try {
		PgqState::Check("create_csr_vertex",
		                pgq_csr_set_degrees(pctx, slot->dev_id, fv.data(), fc.data(), (int64_t)fv.size(), &sum));
	} catch (...) {
		ResetSlot(pctx, *slot); // a failed build leaves nothing behind: the statement can simply be re-run
		throw;
	}
}

// ---- create_csr_edge(INTEGER csr_id, BIGINT v_size, BIGINT num_edges, BIGINT src, BIGINT dst
//                      [, BIGINT edge_rowid [, weight]]) -> INTEGER (R3) -------------------------------------
static void CreateCsrEdgeFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	const idx_t count = args.size();
	result.SetVectorType(VectorType::FLAT_VECTOR);
	if (count == 0) return;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	const int64_t v_size = FirstBigint(args, 1);
	const int64_t num_edges = FirstBigint(args, 2);
	const bool with_eid = args.ColumnCount() >= 6;
	const bool with_w = args.ColumnCount() == 7; // later DuckPGQ: ..., edge_rowid, weight BIGINT|DOUBLE (R10)
	DenseCol src, dst, eid, wcol;
	src.Load(args.data[3], count);
	dst.Load(args.data[4], count);
	if (with_eid) eid.Load(args.data[5], count);
	if (with_w) wcol.Load(args.data[6], count); // BIGINT and DOUBLE are both 8-byte columns: raw copy
	auto out = FlatVector::GetData<int32_t>(result);
	for (idx_t i = 0; i < count; i++) out[i] = 1;
	auto slot = st.Slot(ContextOf(state), csr_id, true);
	PgqExclusiveGuard guard(slot->rw);
	auto *pctx = st.ctx;
	try {
		if (!src.all_valid || !dst.all_valid || (with_eid && !eid.all_valid) || (with_w && !wcol.all_valid)) {
			throw InvalidInputException("pgq_b200: create_csr_edge got a NULL src / dst / edge id / weight");
		}
		if (!slot->vertex_started || slot->finalized) { // edge rows without create_csr_vertex rows
			ResetSlot(pctx, *slot);
			PgqState::Check("create_csr_edge", pgq_csr_create(pctx, slot->dev_id, v_size));
			slot->vertex_started = true;
		}
		if (slot->rows_added == 0) {
			slot->weights.clear();
			slot->wtype = !with_w ? 0
			              : args.data[6].GetType().id() == LogicalTypeId::DOUBLE ? PGQ_WEIGHT_F64 : PGQ_WEIGHT_I64;
		}
		if (with_w) {
			const uint64_t *raw = reinterpret_cast<const uint64_t *>(wcol.val.data());
			slot->weights.insert(slot->weights.end(), raw, raw + count);
		}
		if (!slot->edge_started) { // first row: CsrInitializeEdge
			PgqState::Check("create_csr_edge", pgq_csr_begin_edges(pctx, slot->dev_id, num_edges));
			slot->edge_started = true;
			slot->num_edges = num_edges;
		}
		// rows are appended in arrival order = the stable key of the CSR (deterministic with threads=1)
		PgqState::Check("create_csr_edge", pgq_csr_add_edges(pctx, slot->dev_id, src.val.data(), dst.val.data(),
		                                                     with_eid ? eid.val.data() : nullptr, (int64_t)count));
		slot->rows_added += (int64_t)count;
		if (slot->rows_added == slot->num_edges) FinishSlot(pctx, *slot); // last row
	} catch (...) {
		// an aborted build (NULL row, id out of range, more rows than num_edges, count mismatch ...) leaves
		// nothing behind, so re-running the create_csr_vertex / create_csr_edge statement just works
		ResetSlot(pctx, *slot);
		throw;
	}
}

// ---- cross-call batcher ------------------------------------------------------------------------------------
// One C-ABI call for all queued calls of the leader's kind.  Rows are concatenated in queue order; every call gets
// exactly its own rows' results back (rows are independent: R5 fills lanes in row order and a lane's answer does
// not depend on which other rows share its batch).
static int64_t *PgqReserveVector(void *user, uint64_t total) {
	auto *v = static_cast<std::vector<int64_t> *>(user);
	v->resize(total);
	return v->data();
}

static void RunGroup(pgq_ctx_t *pctx, PgqCsrSlot &slot, std::vector<PgqPending *> &grp) {
	PgqPending &lead = *grp[0];
	const int kind = lead.kind;
	if (grp.size() == 1 && kind != PGQ_K_PATH0 && kind != PGQ_K_PATH1) { // alone: no copies at all
		PgqPending &p = lead;
		int rc = PGQ_OK;
		switch (kind) {
		case PGQ_K_LENGTH:
			rc = pgq_iterativelength(pctx, slot.dev_id, (int64_t)p.n, p.src, p.dst, p.sv, p.dv, (int64_t *)p.out8, p.ov);
			break;
		case PGQ_K_BOUNDED:
			rc = pgq_iterativelength_bounded(pctx, slot.dev_id, (int64_t)p.n, p.src, p.dst, p.sv, p.dv, p.lower, p.upper,
			                                 (int64_t *)p.out8, p.ov);
			break;
		case PGQ_K_REACH:
			rc = pgq_reachability(pctx, slot.dev_id, (int64_t)p.n, p.src, p.dst, p.sv, p.dv, p.outb, p.ov);
			break;
		default:
			rc = pgq_cheapest_path_length(pctx, slot.dev_id, (int64_t)p.n, p.src, p.dst, p.sv, p.dv, p.out8, p.ov);
			break;
		}
		p.rc = rc;
		if (rc != PGQ_OK) p.err = pgq_last_error();
		return;
	}
	uint64_t total = 0;
	for (auto *p : grp) total += p->n;
	std::vector<int64_t> src(total), dst(total);
	std::vector<uint8_t> sv(total), dv(total), ov(total);
	uint64_t at = 0;
	for (auto *p : grp) {
		memcpy(src.data() + at, p->src, p->n * sizeof(int64_t));
		memcpy(dst.data() + at, p->dst, p->n * sizeof(int64_t));
		memcpy(sv.data() + at, p->sv, p->n);
		memcpy(dv.data() + at, p->dv, p->n);
		at += p->n;
	}
	int rc = PGQ_OK;
	std::vector<int64_t> out8;
	std::vector<uint8_t> outb;
	std::vector<pgq_list_entry_t> entries;
	auto vals = std::make_shared<std::vector<int64_t>>();
	switch (kind) {
	case PGQ_K_LENGTH:
		out8.resize(total);
		rc = pgq_iterativelength(pctx, slot.dev_id, (int64_t)total, src.data(), dst.data(), sv.data(), dv.data(),
		                         out8.data(), ov.data());
		break;
	case PGQ_K_BOUNDED:
		out8.resize(total);
		rc = pgq_iterativelength_bounded(pctx, slot.dev_id, (int64_t)total, src.data(), dst.data(), sv.data(), dv.data(),
		                                 lead.lower, lead.upper, out8.data(), ov.data());
		break;
	case PGQ_K_REACH:
		outb.resize(total);
		rc = pgq_reachability(pctx, slot.dev_id, (int64_t)total, src.data(), dst.data(), sv.data(), dv.data(), outb.data(),
		                      ov.data());
		break;
	case PGQ_K_PATH0:
	case PGQ_K_PATH1: {
		entries.resize(total);
		uint64_t nvals = 0;
		rc = pgq_shortestpath_sink(pctx, slot.dev_id, (int64_t)total, src.data(), dst.data(), sv.data(), dv.data(),
		                           kind == PGQ_K_PATH1 ? 1 : 0, entries.data(), ov.data(), PgqReserveVector, vals.get(),
		                           &nvals);
		break;
	}
	default:
		out8.resize(total);
		rc = pgq_cheapest_path_length(pctx, slot.dev_id, (int64_t)total, src.data(), dst.data(), sv.data(), dv.data(),
		                              out8.data(), ov.data());
		break;
	}
	const std::string err = rc != PGQ_OK ? std::string(pgq_last_error()) : std::string();
	at = 0;
	for (auto *p : grp) {
		p->rc = rc;
		p->err = err;
		if (rc == PGQ_OK) {
			memcpy(p->ov, ov.data() + at, p->n);
			if (!out8.empty()) memcpy(p->out8, out8.data() + at, p->n * sizeof(int64_t));
			if (!outb.empty()) memcpy(p->outb, outb.data() + at, p->n);
			if (!entries.empty()) {
				// rows of a call are consecutive, and so are their values: rebase the offsets to the call's slice
				uint64_t first = 0, count = 0;
				bool any = false;
				for (idx_t i = 0; i < p->n; i++) {
					const pgq_list_entry_t &e = entries[at + i];
					if (!ov[at + i]) {
						p->entries[i].offset = any ? count : 0;
						p->entries[i].length = 0;
						continue;
					}
					if (!any) {
						first = e.offset;
						any = true;
					}
					p->entries[i].offset = e.offset - first;
					p->entries[i].length = e.length;
					count = e.offset - first + e.length;
				}
				p->vals_all = vals;
				p->vals_begin = first;
				p->vals_count = count;
			}
		}
		at += p->n;
	}
}

// Called with the slot's shared lock held.  Returns when p is done (by this thread as leader, or by another).
static void SubmitQuery(pgq_ctx_t *pctx, PgqCsrSlot &slot, PgqPending &p) {
	std::unique_lock<std::mutex> l(slot.bm);
	slot.queue.push_back(&p);
	slot.calls++;
	while (!p.done) {
		if (slot.leader_active) {
			slot.bcv.wait(l);
			continue;
		}
		// become the leader: take every queued call of my kind (mine is among them), leave the rest queued
		slot.leader_active = true;
		std::vector<PgqPending *> grp, rest;
		for (auto *q : slot.queue) (q->SameGroup(p) ? grp : rest).push_back(q);
		slot.queue.swap(rest);
		slot.gpu_calls++;
		l.unlock();
		try {
			RunGroup(pctx, slot, grp);
		} catch (std::exception &ex) { // allocation failure while concatenating: everybody in the group hears about it
			for (auto *q : grp) {
				q->rc = PGQ_ERR_OOM;
				q->err = ex.what();
			}
		}
		l.lock();
		for (auto *q : grp) q->done = true;
		slot.leader_active = false;
		slot.bcv.notify_all();
	}
}

static void ThrowIfFailed(const char *what, const PgqPending &p) {
	if (p.rc != PGQ_OK) PgqState::Fail(what, p.rc, p.err.c_str());
}

// ---- iterativelength / reachability (R4-R6) ----------------------------------------------------------
template <bool REACH>
static void LengthFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	idx_t count = args.size();
	const bool constant = AllConstant(args);
	result.SetVectorType(constant ? VectorType::CONSTANT_VECTOR : VectorType::FLAT_VECTOR);
	if (count == 0) return;
	if (constant) count = 1;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	// reachability's optional BOOLEAN variant flag sits at position 1 and does not change results
	const idx_t base = args.ColumnCount() == 5 ? 2 : 1;
	DenseCol src, dst;
	src.Load(args.data[base + 1], count);
	dst.Load(args.data[base + 2], count);
	std::vector<uint8_t> ov(count), ob;
	PgqPending p;
	p.n = count;
	p.src = src.val.data();
	p.dst = dst.val.data();
	p.sv = src.valid.data();
	p.dv = dst.valid.data();
	p.ov = ov.data();
	if (REACH) {
		ob.resize(count);
		p.kind = PGQ_K_REACH;
		p.outb = ob.data();
	} else {
		p.kind = PGQ_K_LENGTH;
		p.out8 = constant ? (void *)ConstantVector::GetData<int64_t>(result) : (void *)FlatVector::GetData<int64_t>(result);
		if (args.ColumnCount() == 6) {
			// iterativelength(csr_id, v_size, src, dst, lower, upper) of the later DuckPGQ code (R11): the
			// bounds are per-query constants, read from row 0 like csr_id / v_size
			p.kind = PGQ_K_BOUNDED;
			p.lower = FirstBigint(args, 4);
			p.upper = FirstBigint(args, 5);
			if (p.lower < 0 || p.upper < 0)
				throw InvalidInputException("pgq_b200: iterativelength bounds must be non-negative");
		}
	}
	auto slot = ReadySlot(ContextOf(state), csr_id);
	{
		PgqSharedGuard guard(slot->rw);
		if (slot->dropped || !slot->finalized)
			throw InvalidInputException("pgq_b200: csr " + std::to_string(csr_id) + " was deleted while in use");
		SubmitQuery(st.ctx, *slot, p);
	}
	ThrowIfFailed(REACH ? "reachability" : "iterativelength", p);
	if (constant) {
		if (REACH) ConstantVector::GetData<bool>(result)[0] = ob[0] != 0;
		ConstantVector::SetNull(result, !ov[0]);
		return;
	}
	auto &validity = FlatVector::Validity(result);
	if (REACH) {
		auto out = FlatVector::GetData<bool>(result);
		for (idx_t i = 0; i < count; i++) out[i] = ob[i] != 0;
	}
	for (idx_t i = 0; i < count; i++) {
		if (!ov[i]) validity.SetInvalid(i);
	}
}

// ---- cheapest_path_length -> BIGINT / DOUBLE (R10) --------------------------------------------------
// The result type of a scalar function is fixed at bind time, before the CSR (and its weight type)
// exists, so the two weight types get two names: cheapest_path_length (BIGINT weights) and
// cheapest_path_length_double (DOUBLE weights).
template <bool F64>
static void CheapestPathLengthFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	idx_t count = args.size();
	const bool constant = AllConstant(args);
	result.SetVectorType(constant ? VectorType::CONSTANT_VECTOR : VectorType::FLAT_VECTOR);
	if (count == 0) return;
	if (constant) count = 1;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	DenseCol src, dst;
	src.Load(args.data[2], count);
	dst.Load(args.data[3], count);
	std::vector<uint8_t> ov(count);
	auto slot = ReadySlot(ContextOf(state), csr_id);
	PgqPending p;
	p.kind = F64 ? PGQ_K_CHEAP_F64 : PGQ_K_CHEAP_I64;
	p.n = count;
	p.src = src.val.data();
	p.dst = dst.val.data();
	p.sv = src.valid.data();
	p.dv = dst.valid.data();
	p.ov = ov.data();
	if (constant)
		p.out8 = F64 ? (void *)ConstantVector::GetData<double>(result) : (void *)ConstantVector::GetData<int64_t>(result);
	else p.out8 = F64 ? (void *)FlatVector::GetData<double>(result) : (void *)FlatVector::GetData<int64_t>(result);
	{
		PgqSharedGuard guard(slot->rw);
		if (slot->dropped || !slot->finalized)
			throw InvalidInputException("pgq_b200: csr " + std::to_string(csr_id) + " was deleted while in use");
		int wtype = 0;
		PgqState::Check("cheapest_path_length", pgq_csr_weight_type(st.ctx, slot->dev_id, &wtype));
		if (wtype != (F64 ? PGQ_WEIGHT_F64 : PGQ_WEIGHT_I64)) {
			throw InvalidInputException(wtype == 0 ? "pgq_b200: this CSR has no weights (7-argument create_csr_edge)"
			                            : F64      ? "pgq_b200: BIGINT weights: use cheapest_path_length"
			                                       : "pgq_b200: DOUBLE weights: use cheapest_path_length_double");
		}
		SubmitQuery(st.ctx, *slot, p);
	}
	ThrowIfFailed("cheapest_path_length", p);
	if (constant) {
		ConstantVector::SetNull(result, !ov[0]);
		return;
	}
	auto &validity = FlatVector::Validity(result);
	for (idx_t i = 0; i < count; i++) {
		if (!ov[i]) validity.SetInvalid(i);
	}
}

// ---- shortestpath -> LIST<BIGINT> (R7-R9) -------------------------------------------------------------
template <int MODE>
static void ShortestPathFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	idx_t count = args.size();
	const bool constant = AllConstant(args);
	result.SetVectorType(constant ? VectorType::CONSTANT_VECTOR : VectorType::FLAT_VECTOR);
	if (count == 0) return;
	if (constant) count = 1;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	DenseCol src, dst;
	src.Load(args.data[2], count);
	dst.Load(args.data[3], count);
	std::vector<uint8_t> ov(count);
	std::vector<pgq_list_entry_t> entries(count);
	PgqPending p;
	p.kind = MODE ? PGQ_K_PATH1 : PGQ_K_PATH0;
	p.n = count;
	p.src = src.val.data();
	p.dst = dst.val.data();
	p.sv = src.valid.data();
	p.dv = dst.valid.data();
	p.ov = ov.data();
	p.entries = entries.data();
	auto slot = ReadySlot(ContextOf(state), csr_id);
	{
		PgqSharedGuard guard(slot->rw);
		if (slot->dropped || !slot->finalized)
			throw InvalidInputException("pgq_b200: csr " + std::to_string(csr_id) + " was deleted while in use");
		SubmitQuery(st.ctx, *slot, p);
	}
	ThrowIfFailed("shortestpath", p);
	// the list child vector is filled straight from the batch's value buffer (one copy, no malloc)
	auto list_data = constant ? ConstantVector::GetData<list_entry_t>(result) : FlatVector::GetData<list_entry_t>(result);
	for (idx_t i = 0; i < count; i++) {
		list_data[i].offset = entries[i].offset;
		list_data[i].length = ov[i] ? entries[i].length : 0;
	}
	const uint64_t total = p.vals_count;
	ListVector::Reserve(result, total);
	if (total)
		memcpy(FlatVector::GetData<int64_t>(ListVector::GetEntry(result)), p.vals_all->data() + p.vals_begin,
		       total * sizeof(int64_t));
	ListVector::SetListSize(result, total);
	if (constant) {
		ConstantVector::SetNull(result, !ov[0]);
		return;
	}
	auto &validity = FlatVector::Validity(result);
	for (idx_t i = 0; i < count; i++) {
		if (!ov[i]) validity.SetInvalid(i);
	}
}

// ---- delete_csr(INTEGER csr_id) -> BOOLEAN ------------------------------------------------------------
static void DeleteCsrFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	const idx_t count = args.size();
	result.SetVectorType(VectorType::FLAT_VECTOR);
	if (count == 0) return;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	auto out = FlatVector::GetData<bool>(result);
	auto slot = st.Slot(ContextOf(state), csr_id, false);
	const bool existed = slot != nullptr;
	if (existed) {
		{
			PgqExclusiveGuard guard(slot->rw); // waits for the queries in flight on this CSR
			ResetSlot(st.ctx, *slot);
			slot->dropped = true;
		}
		std::lock_guard<std::mutex> lock(st.mu);
		st.slots.erase(std::make_pair(ContextOf(state), csr_id));
	}
	for (idx_t i = 0; i < count; i++) out[i] = existed;
}

// ---- pgq_batcher_stats(INTEGER csr_id) -> VARCHAR: query calls entered / C-ABI calls issued for them --------
static void BatcherStatsFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	const idx_t count = args.size();
	result.SetVectorType(VectorType::FLAT_VECTOR);
	if (count == 0) return;
	auto slot = State().Slot(ContextOf(state), FirstInt(args, 0), false);
	string text = "calls=0 gpu_calls=0";
	if (slot) {
		std::lock_guard<std::mutex> l(slot->bm);
		text = "calls=" + std::to_string(slot->calls) + " gpu_calls=" + std::to_string(slot->gpu_calls);
	}
	auto out = FlatVector::GetData<string_t>(result);
	for (idx_t i = 0; i < count; i++) out[i] = StringVector::AddString(result, text);
}

// ---- pgq_rewrite(VARCHAR statement) -> VARCHAR: the SQL/PGQ rewriter (pgq_rewrite.hpp) for clients that cannot put it
// in front of the parser: CREATE / DROP PROPERTY GRAPH register in a process-wide catalog and return '', a statement
// with GRAPH_TABLE comes back as plain SQL (the query first, then one "SELECT delete_csr(id);" per CSR it builds, one
// statement per line), anything else comes back unchanged.  The client runs what it gets.  Errors raise. -------------
static void RewriteFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	static std::mutex mu;
	static pgq_rewrite::Rewriter rewriter;
	const idx_t count = args.size();
	result.SetVectorType(VectorType::FLAT_VECTOR);
	VectorData vd;
	args.data[0].Orrify(count, vd);
	auto in = (const string_t *)vd.data;
	auto out = FlatVector::GetData<string_t>(result);
	for (idx_t i = 0; i < count; i++) {
		const idx_t k = vd.sel->get_index(i);
		if (!vd.validity.RowIsValid(k)) {
			FlatVector::Validity(result).SetInvalid(i);
			continue;
		}
		const string text = in[k].GetString();
		pgq_rewrite::Result r;
		{
			std::lock_guard<std::mutex> l(mu);
			r = rewriter.rewrite(text);
		}
		if (!r.error.empty()) throw InvalidInputException(r.error);
		string res;
		if (!r.handled) res = text;
		else
			for (auto &q : r.statements) {
				string one = q;
				for (auto &ch : one)
					if (ch == '\n') ch = ' ';
				res += (res.empty() ? "" : "\n") + one;
			}
		out[i] = StringVector::AddString(result, res);
	}
}

static void Register(ClientContext &context, Catalog &catalog, const string &name, vector<LogicalType> args,
                     LogicalType ret, scalar_function_t fn) {
	ScalarFunction f(name, move(args), move(ret), move(fn), /*has_side_effects=*/true, PgqBind);
	CreateScalarFunctionInfo info(f);
	info.on_conflict = OnCreateConflict::ERROR_ON_CONFLICT;
	catalog.CreateFunction(context, &info);
}

} // namespace duckdb

using namespace duckdb;

extern "C" {
DUCKDB_EXTENSION_API void pgq_b200_init(duckdb::DatabaseInstance &db) {
	Connection con(db);
	con.BeginTransaction();
	auto &context = *con.context;
	auto &catalog = Catalog::GetCatalog(context);
	const LogicalType I = LogicalType::INTEGER, B = LogicalType::BIGINT, BO = LogicalType::BOOLEAN;
	Register(context, catalog, "create_csr_vertex", {I, B, B, B}, B, CreateCsrVertexFunction);
	{
		// two overloads: with and without the edge rowid column
		ScalarFunctionSet set("create_csr_edge");
		set.AddFunction(ScalarFunction({I, B, B, B, B}, I, CreateCsrEdgeFunction, true, PgqBind));
		set.AddFunction(ScalarFunction({I, B, B, B, B, B}, I, CreateCsrEdgeFunction, true, PgqBind));
		// with an edge weight (later DuckPGQ, rule R10)
		set.AddFunction(ScalarFunction({I, B, B, B, B, B, B}, I, CreateCsrEdgeFunction, true, PgqBind));
		set.AddFunction(ScalarFunction({I, B, B, B, B, B, LogicalType::DOUBLE}, I, CreateCsrEdgeFunction, true,
		                               PgqBind));
		CreateScalarFunctionInfo info(set);
		catalog.CreateFunction(context, &info);
	}
	{
		// 4 arguments, and the later 6-argument form with lower / upper length bounds (f1, rule R11)
		ScalarFunctionSet set("iterativelength");
		set.AddFunction(ScalarFunction({I, B, B, B}, B, LengthFunction<false>, true, PgqBind));
		set.AddFunction(ScalarFunction({I, B, B, B, B, B}, B, LengthFunction<false>, true, PgqBind));
		CreateScalarFunctionInfo info(set);
		catalog.CreateFunction(context, &info);
	}
	{
		ScalarFunctionSet set("reachability");
		set.AddFunction(ScalarFunction({I, B, B, B}, BO, LengthFunction<true>, true, PgqBind));
		set.AddFunction(ScalarFunction({I, BO, B, B, B}, BO, LengthFunction<true>, true, PgqBind));
		CreateScalarFunctionInfo info(set);
		catalog.CreateFunction(context, &info);
	}
	Register(context, catalog, "shortestpath", {I, B, B, B}, LogicalType::LIST(B), ShortestPathFunction<0>);
	// later DuckPGQ behaviour: [v, e, v, ..., v] with edge row ids interleaved
	Register(context, catalog, "shortestpath_with_edges", {I, B, B, B}, LogicalType::LIST(B),
	         ShortestPathFunction<1>);
	Register(context, catalog, "cheapest_path_length", {I, B, B, B}, B, CheapestPathLengthFunction<false>);
	Register(context, catalog, "cheapest_path_length_double", {I, B, B, B}, LogicalType::DOUBLE,
	         CheapestPathLengthFunction<true>);
	Register(context, catalog, "delete_csr", {I}, BO, DeleteCsrFunction);
	Register(context, catalog, "pgq_batcher_stats", {I}, LogicalType::VARCHAR, BatcherStatsFunction);
	Register(context, catalog, "pgq_rewrite", {LogicalType::VARCHAR}, LogicalType::VARCHAR, RewriteFunction);
	con.Commit();
}

DUCKDB_EXTENSION_API const char *pgq_b200_version() {
	return DuckDB::LibraryVersion();
}
}
