// Package gpu_plugin: the same kube-scheduler plugin surface as the reference
// (pkg/plugins/gpu_plugin/gpu_plugins.go: Name :50, New :928, Score :779, ScoreExtensions
// :812, NormalizeScore :816, PostBind :843) with the per-(pod,node) arithmetic moved to the
// GPU.  PreScore (allowed by the framework API) scores the cycle's pod against every node
// in one call and stashes the row in CycleState; Score becomes a lookup; NormalizeScore
// keeps the reference's contract on the framework's own NodeScoreList.
//
// NOT BUILT in the development container (no Go toolchain, k8s.io/kubernetes v1.21.0 not
// vendored).  The snapshot builder (informer caches + one Redis MGET + one bulk recommender
// call -> gpusched.Cluster / gpusched.Predictions) is SURVEY 8(f) item 1 and is sketched in
// INTEGRATION.md.
package gpu_plugin

import (
	"context"
	"fmt"
	"strconv"
	"sync"

	corev1 "k8s.io/api/core/v1"
	"k8s.io/apimachinery/pkg/runtime"
	"k8s.io/kubernetes/pkg/scheduler/framework"

	"github.com/dimgatz98/k8s-gpu-scheduler/pkg/gpusched"
	"github.com/dimgatz98/k8s-gpu-scheduler/pkg/utils"
)

const Name = "GPU"
const stateKey = "GPU/row"

type GPU struct {
	handle framework.Handle
	mu     sync.RWMutex // guards snap (rebuilt on informer / Redis / recommender events)
	ctx    *gpusched.Ctx
	snap   *Snapshot
}

// Snapshot maps the string world of the reference onto table indices.
type Snapshot struct {
	NodeIndex map[string]int    // nodeName -> row of the node tables
	UUIDs     [][]string        // per node, Redis array order
	AppOf     func(string) uint32 // pod name -> app id; first table key contained in the name
}

var _ = framework.PreScorePlugin(&GPU{})
var _ = framework.ScorePlugin(&GPU{})
var _ = framework.PostBindPlugin(&GPU{})

type row struct {
	score []uint8 // raw Score() per node index
	sel   []uint8 // packed selection byte per node index
}

func (r *row) Clone() framework.StateData { return r }

func (g *GPU) Name() string { return Name }

// currSLO, gpu_plugins.go:459-469
func podSLO(pod *corev1.Pod) float32 {
	v := utils.GetEnv(pod, "SLO")
	if v == "" {
		return 0
	}
	f, err := strconv.ParseFloat(v, 32)
	if err != nil {
		return 0
	}
	return float32(f)
}

func (g *GPU) PreScore(ctx context.Context, state *framework.CycleState, pod *corev1.Pod, nodes []*corev1.Node) *framework.Status {
	g.mu.RLock()
	defer g.mu.RUnlock()
	n := len(g.snap.NodeIndex)
	r := &row{score: make([]uint8, n), sel: make([]uint8, n)}
	app := []uint32{g.snap.AppOf(pod.Name)}
	slo := []float32{podSLO(pod)}
	if err := g.ctx.ScoreBatch(app, slo, false, r.score, r.sel, nil); err != nil {
		return framework.NewStatus(framework.Error, fmt.Sprintf("Error in Logic() in Score(): %s", err))
	}
	state.Write(stateKey, r)
	return nil
}

func (g *GPU) Score(ctx context.Context, state *framework.CycleState, pod *corev1.Pod, nodeName string) (int64, *framework.Status) {
	d, err := state.Read(stateKey)
	if err != nil {
		return 0, framework.NewStatus(framework.Error, fmt.Sprintf("Error in Logic() in Score(): %s", err))
	}
	g.mu.RLock()
	i, ok := g.snap.NodeIndex[nodeName]
	g.mu.RUnlock()
	if !ok {
		return 0, framework.NewStatus(framework.Error, fmt.Sprintf("getting node %q from Snapshot: not found", nodeName))
	}
	// the selected UUID (gpu_plugins.go:760-768) is snap.UUIDs[i][sel&31] when sel&31 != 31;
	// PostBind writes it to the pod's ConfigMap exactly as the reference does (:907-920)
	return int64(d.(*row).score[i]), nil
}

func (g *GPU) ScoreExtensions() framework.ScoreExtensions { return g }

// NormalizeScore: gpu_plugins.go:816-841 on the framework's list (feasible nodes only).
func (g *GPU) NormalizeScore(ctx context.Context, state *framework.CycleState, pod *corev1.Pod, scores framework.NodeScoreList) *framework.Status {
	tmp := make([]int64, len(scores))
	for i := range scores {
		tmp[i] = scores[i].Score
	}
	if err := g.ctx.NormalizeScores(tmp); err != nil {
		return nil // the reference's NormalizeScore always returns nil (:840)
	}
	for i := range scores {
		scores[i].Score = tmp[i]
	}
	return nil
}

func (g *GPU) PostBind(ctx context.Context, state *framework.CycleState, p *corev1.Pod, nodeName string) {
	// unchanged control-plane work of the reference (gpu_plugins.go:843-926): read the selected
	// UUID for nodeName, translate MPS_<node>, write CUDA_VISIBLE_DEVICES etc. to the ConfigMap.
}

func New(_ runtime.Object, handle framework.Handle) (framework.Plugin, error) {
	c, err := gpusched.New(0)
	if err != nil {
		return nil, err
	}
	return &GPU{handle: handle, ctx: c}, nil
}
