// TEST INFRASTRUCTURE ONLY -- a stand-in for libbt2gpu.so that answers bt2gpu_create / bt2gpu_align_batch with the ORACLE
// (oracle_capi.cpp), so that the reference-side binding (integration/gpu_batch.cpp + bt2_search_gpu.patch, built into
// oracle/_ref/bowtie2-align-l-gpu) can be exercised on a machine without a GPU: LD_PRELOAD=tests/emu/libbt2gpu_shim.so.
// It checks the BINDING (SoA batch out of the reference's slots, replay into AlnSinkWrap, the aliased edit buffer), not the
// CUDA path; tests/test_gpu_dropin.py runs the same binary against the real library.  Never shipped, never loaded by the product.
#include <cstring>
#include <string>
#include <vector>
#include "bt2gpu.h"

extern "C" {
struct oracle_ctx;
oracle_ctx* oracle_create(const bt2gpu_index*, const bt2gpu_ref*, const bt2gpu_params*);
void oracle_destroy(oracle_ctx*);
int oracle_align_batch(oracle_ctx*, const bt2gpu_reads*, bt2gpu_results*, int);
}

struct bt2gpu_ctx {
	bt2gpu_index ix; bt2gpu_ref rf; bt2gpu_params p; oracle_ctx* o = nullptr; std::string err;
	std::vector<uint64_t> plen, rstarts, fchr, ftab, eftab, offs, roff, rlen; std::vector<uint8_t> ebwt, rfirst, buf;
};

extern "C" {
int bt2gpu_abi_version(void) { return BT2GPU_ABI_VERSION; }
void* bt2gpu_host_alloc(size_t bytes) { return malloc(bytes ? bytes : 1); }
void bt2gpu_host_free(void* p) { free(p); }
int bt2gpu_max_rows(void) { return BT2GPU_MAX_ROWS; }
int bt2gpu_max_cols(void) { return BT2GPU_MAX_COLS; }
bt2gpu_ctx* bt2gpu_create(const bt2gpu_index* ix, const bt2gpu_ref* rf, const bt2gpu_params* p, int, char* err, size_t errlen) {
	if(p->abi_version != BT2GPU_ABI_VERSION) { if(err) snprintf(err, errlen, "shim: ABI mismatch"); return nullptr; }
	bt2gpu_ctx* c = new bt2gpu_ctx();      // deep copies: like the real library, nothing of the caller's is kept
	c->ix = *ix; c->rf = *rf; c->p = *p;
	c->plen.assign(ix->plen, ix->plen + ix->nPat); c->rstarts.assign(ix->rstarts, ix->rstarts + 3 * ix->nFrag);
	c->fchr.assign(ix->fchr, ix->fchr + 5); c->ftab.assign(ix->ftab, ix->ftab + ix->ftabLen); c->eftab.assign(ix->eftab, ix->eftab + ix->eftabLen);
	c->offs.assign(ix->offs, ix->offs + ix->offsLen); c->ebwt.assign(ix->ebwt, ix->ebwt + ix->ebwtTotLen);
	c->roff.assign(rf->rec_off, rf->rec_off + rf->nrecs); c->rlen.assign(rf->rec_len, rf->rec_len + rf->nrecs);
	c->rfirst.assign(rf->rec_first, rf->rec_first + rf->nrecs); c->buf.assign(rf->buf, rf->buf + (rf->bufBases + 3) / 4);
	c->ix.plen = c->plen.data(); c->ix.rstarts = c->rstarts.data(); c->ix.fchr = c->fchr.data(); c->ix.ftab = c->ftab.data();
	c->ix.eftab = c->eftab.data(); c->ix.offs = c->offs.data(); c->ix.ebwt = c->ebwt.data();
	c->rf.rec_off = c->roff.data(); c->rf.rec_len = c->rlen.data(); c->rf.rec_first = c->rfirst.data(); c->rf.buf = c->buf.data();
	c->o = oracle_create(&c->ix, &c->rf, &c->p);
	return c;
}
int bt2gpu_align_batch(bt2gpu_ctx* c, const bt2gpu_reads* in, bt2gpu_results* out) { return oracle_align_batch(c->o, in, out, 1); }
const char* bt2gpu_last_error(const bt2gpu_ctx* c) { return c ? c->err.c_str() : "null"; }
void bt2gpu_destroy(bt2gpu_ctx* c) { if(c) { oracle_destroy(c->o); delete c; } }
}
