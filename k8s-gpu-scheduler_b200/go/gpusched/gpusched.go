// Package gpusched is the cgo binding of libgpusched.so (include/gpusched.h).
//
// It is the only place the Go scheduler touches C.  No Go pointer is retained by C after a
// call returns (the library copies what it needs): flat arguments are passed as plain slices,
// table structs are filled with C copies of the slices (cgo forbids Go pointers inside C memory).
// unsafe.Slice needs Go 1.17+; the reference's go.mod says 1.16 — bump it or copy in a loop.  NOT BUILT in the development container (no Go toolchain there); kept next
// to the library so that a maintainer of dimgatz98/k8s-gpu-scheduler can vendor it under
// pkg/gpusched.
package gpusched

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../.. -lgpusched -Wl,-rpath,${SRCDIR}/../..
#include <stdlib.h>
#include "gpusched.h"
*/
import "C"

import (
	"fmt"
	"unsafe"
)

const (
	ModelNone  = C.GS_MODEL_NONE
	ModelA30   = C.GS_MODEL_A30
	ModelV100  = C.GS_MODEL_V100
	AppUnknown = uint32(C.GS_APP_UNKNOWN)
	SelNone    = 31
)

// Cluster is the structure-of-arrays snapshot that replaces the per-Score Redis GET
// (gpu_plugins.go:529-542) and GetSLOs (gpu_plugins.go:87-160).
type Cluster struct {
	MaxGPUs        int
	NodeModel      []uint8   // [N]
	NodeNGPU       []uint8   // [N]
	GPUOff         []uint32  // [N*MaxGPUs+1]
	ResApp         []uint32  // [R]
	ResSLO         []float32 // [R]
	RedisReachable bool
}

// Predictions are the completed recommender tables (gpu_plugins.go:302-355).
type Predictions struct {
	NApps, NCols int
	Conf         []float32 // [NApps*NCols]
	Interf       []float32 // [2*NApps*NApps]
	ColOf        []int32   // [2*32]
}

type Ctx struct{ h *C.gs_ctx }

func (c *Ctx) err(rc C.int) error {
	if rc == 0 {
		return nil
	}
	return fmt.Errorf("gpusched: %d: %s", int(rc), C.GoString(C.gs_last_error(c.h)))
}

func New(device int) (*Ctx, error) {
	var h *C.gs_ctx
	if rc := C.gs_create(C.int(device), &h); rc != 0 {
		return nil, fmt.Errorf("gpusched: gs_create(%d) = %d (no CUDA device? there is no CPU fallback)", device, int(rc))
	}
	return &Ctx{h: h}, nil
}

func (c *Ctx) Close() { C.gs_destroy(c.h); c.h = nil }

func p8(s []uint8) *C.uint8_t {
	if len(s) == 0 {
		return nil
	}
	return (*C.uint8_t)(unsafe.Pointer(&s[0]))
}
func p32(s []uint32) *C.uint32_t {
	if len(s) == 0 {
		return nil
	}
	return (*C.uint32_t)(unsafe.Pointer(&s[0]))
}
func pf(s []float32) *C.float {
	if len(s) == 0 {
		return nil
	}
	return (*C.float)(unsafe.Pointer(&s[0]))
}

// cbuf copies a Go slice into C memory: the table structs hold pointers, and cgo forbids storing
// Go pointers in C memory (the library copies everything to the device before it returns, so the
// C copies are freed right after the call).
func cbuf(p unsafe.Pointer, n int) unsafe.Pointer {
	if n == 0 {
		return nil
	}
	return C.CBytes(unsafe.Slice((*byte)(p), n))
}

func (c *Ctx) UploadCluster(cl *Cluster, nodeBase int) error {
	var cc C.gs_cluster
	cc.n_nodes = C.int32_t(len(cl.NodeModel))
	cc.max_gpus = C.int32_t(cl.MaxGPUs)
	bufs := []unsafe.Pointer{
		cbuf(unsafe.Pointer(p8(cl.NodeModel)), len(cl.NodeModel)),
		cbuf(unsafe.Pointer(p8(cl.NodeNGPU)), len(cl.NodeNGPU)),
		cbuf(unsafe.Pointer(p32(cl.GPUOff)), 4*len(cl.GPUOff)),
		cbuf(unsafe.Pointer(p32(cl.ResApp)), 4*len(cl.ResApp)),
		cbuf(unsafe.Pointer(pf(cl.ResSLO)), 4*len(cl.ResSLO)),
	}
	defer func() {
		for _, b := range bufs {
			C.free(b)
		}
	}()
	cc.node_model, cc.node_ngpu = (*C.uint8_t)(bufs[0]), (*C.uint8_t)(bufs[1])
	cc.gpu_off, cc.res_app, cc.res_slo = (*C.uint32_t)(bufs[2]), (*C.uint32_t)(bufs[3]), (*C.float)(bufs[4])
	if cl.RedisReachable {
		cc.redis_reachable = 1
	}
	return c.err(C.gs_upload_cluster(c.h, &cc, C.int32_t(nodeBase)))
}

func (c *Ctx) UploadPredictions(p *Predictions) error {
	var cp C.gs_predictions
	cp.n_apps, cp.n_cols = C.int32_t(p.NApps), C.int32_t(p.NCols)
	bufs := []unsafe.Pointer{
		cbuf(unsafe.Pointer(pf(p.Conf)), 4*len(p.Conf)),
		cbuf(unsafe.Pointer(pf(p.Interf)), 4*len(p.Interf)),
		cbuf(unsafe.Pointer(&p.ColOf[0]), 4*len(p.ColOf)),
	}
	defer func() {
		for _, b := range bufs {
			C.free(b)
		}
	}()
	cp.conf, cp.interf, cp.col_of = (*C.float)(bufs[0]), (*C.float)(bufs[1]), (*C.int32_t)(bufs[2])
	return c.err(C.gs_upload_predictions(c.h, &cp))
}

// ScoreBatch = Score -> Logic for every (pod,node) and NormalizeScore per pod
// (gpu_plugins.go:779-810, 455-777, 816-841).  score/sel are [len(app)*N]; best is [len(app)].
func (c *Ctx) ScoreBatch(app []uint32, slo []float32, normalize bool, score, sel []uint8, best []int32) error {
	flags := C.uint32_t(0)
	if normalize {
		flags = C.GS_SCORE_NORMALIZE
	}
	var pb *C.int32_t
	if len(best) > 0 {
		pb = (*C.int32_t)(unsafe.Pointer(&best[0]))
	}
	return c.err(C.gs_score_batch(c.h, p32(app), pf(slo), nil, C.int32_t(len(app)), flags, p8(score), p8(sel), pb))
}

// NormalizeScores = NormalizeScore on one NodeScoreList (gpu_plugins.go:816-841).
func (c *Ctx) NormalizeScores(scores []int64) error {
	if len(scores) == 0 {
		return nil
	}
	return c.err(C.gs_normalize_scores(c.h, (*C.int64_t)(unsafe.Pointer(&scores[0])), C.int32_t(len(scores))))
}

// SelectMIG = the configuration choice of reconfigure() (gpu_plugins.go:366-399).
func (c *Ctx) SelectMIG(app []uint32, slo []float32, out []uint8) error {
	return c.err(C.gs_select_mig(c.h, p32(app), pf(slo), C.int32_t(len(app)), p8(out)))
}
