package routing

// UNVERIFIED SOURCE (no Go toolchain where this was written) — a transliteration of
// polykey_service_b200/host/batching_service.cpp, which is compiled and tested.

import (
	"context"
	"errors"
	"fmt"
	"sync"
	"sync/atomic"
	"time"
	"unsafe"

	"github.com/spounge-ai/polykey-service/internal/service"
	cmn "github.com/spounge-ai/spounge-proto/gen/go/common/v2"
	pk "github.com/spounge-ai/spounge-proto/gen/go/polykey/v2"
	"google.golang.org/protobuf/encoding/protojson"
	"google.golang.org/protobuf/types/known/structpb"
)

// Config: the knobs follow the reference's pattern of env vars read in main
// (cmd/polykey/main.go:57-60): PK_MAX_BATCH, PK_MAX_WAIT_US, PK_IDLE_FLUSH_US, PK_GPUS.
type Config struct {
	MaxBatch        int           // requests per flush
	MaxWait         time.Duration // deadline from the first queued request
	IdleFlush       time.Duration // flush earlier once nothing has arrived for this long
	// HoldBatches switches the work-conserving default off.  By default a free slot takes whatever
	// is queued right now and batches grow by themselves while all slots are busy — measured with
	// the C++ mirror (profiles/README.md §5): p50 25 µs instead of 105 µs at 1e5 requests/s, because
	// with arrivals closer than IdleFlush apart the idle window almost never closes.
	HoldBatches bool
	MaxPayloadBytes int           // arena bytes per batch
	Devices         []int         // one flusher, one Router, one table replica per GPU
	Slots           int           // batches in flight per GPU
}

type pending struct {
	name     string
	payload  []byte
	decision Decision
	err      error
	done     chan struct{}
}

type arena struct {
	payload []byte
	payOff  []uint64
	payLen  []uint32
	names   []byte
	nameOff []uint32
	out     []Decision
	reqs    []*pending
}

type shard struct {
	r      *Router
	queue  chan *pending
	arenas []arena
}

// BatchingService implements service.Service (internal/service/service.go:13-15).  It is
// handed to server.NewServer where NewMockService() is today (cmd/polykey/main.go:85-86).
type BatchingService struct {
	inner   service.Service
	cfg     Config
	shards  []*shard
	nextID  atomic.Uint64
	healthy atomic.Bool
	stop    chan struct{}
	wg      sync.WaitGroup
}

var _ service.Service = (*BatchingService)(nil)

type decisionKey struct{}

// DecisionFromContext lets the inner service (or an interceptor) read the routing decision;
// the proto surface is untouched.
func DecisionFromContext(ctx context.Context) (Decision, bool) {
	d, ok := ctx.Value(decisionKey{}).(Decision)
	return d, ok
}

func NewBatchingService(inner service.Service, cfg Config) (*BatchingService, error) {
	b := &BatchingService{inner: inner, cfg: cfg, stop: make(chan struct{})}
	for _, dev := range cfg.Devices {
		r, err := NewRouter(dev, uint32(cfg.MaxBatch), uint64(cfg.MaxPayloadBytes), uint32(cfg.Slots))
		if err != nil {
			b.Shutdown()
			return nil, err
		}
		s := &shard{r: r, queue: make(chan *pending, 4*cfg.MaxBatch), arenas: make([]arena, cfg.Slots)}
		for i := range s.arenas {
			a := &s.arenas[i]
			a.payload = r.HostAlloc(cfg.MaxPayloadBytes + 64)
			a.names = r.HostAlloc(256 * cfg.MaxBatch)
			a.payOff = unsafe.Slice((*uint64)(unsafe.Pointer(unsafe.SliceData(r.HostAlloc(8*(cfg.MaxBatch+1))))), cfg.MaxBatch+1)
			a.payLen = unsafe.Slice((*uint32)(unsafe.Pointer(unsafe.SliceData(r.HostAlloc(4*cfg.MaxBatch)))), cfg.MaxBatch)
			a.nameOff = unsafe.Slice((*uint32)(unsafe.Pointer(unsafe.SliceData(r.HostAlloc(4*(cfg.MaxBatch+1))))), cfg.MaxBatch+1)
			a.out = unsafe.Slice((*Decision)(unsafe.Pointer(unsafe.SliceData(r.HostAlloc(24*cfg.MaxBatch)))), cfg.MaxBatch)
		}
		b.shards = append(b.shards, s)
		b.wg.Add(1)
		go b.flusher(s)
	}
	if len(b.shards) == 0 {
		return nil, errors.New("routing: no devices configured")
	}
	b.healthy.Store(true)
	return b, nil
}

// Healthy turns false after a device error: main should set health to NOT_SERVING
// (cmd/polykey/main.go:93-94).
func (b *BatchingService) Healthy() bool { return b.healthy.Load() }

// Shutdown drains the queues and releases the GPUs; call it before srv.GracefulStop()
// (cmd/polykey/main.go:118-119).
func (b *BatchingService) Shutdown() {
	select {
	case <-b.stop:
	default:
		close(b.stop)
	}
	b.wg.Wait()
	for _, s := range b.shards {
		s.r.Close()
	}
}

// ExecuteTool: enqueue, wait for the batch this request landed in, then let the inner service
// build the response exactly as before (internal/service/mock.go:22-67).
func (b *BatchingService) ExecuteTool(ctx context.Context, toolName string, parameters *structpb.Struct,
	secretId *string, metadata *cmn.Metadata) (*pk.ExecuteToolResponse, error) {
	if !b.healthy.Load() {
		return nil, errors.New("routing stage is unhealthy")
	}
	p := &pending{name: toolName, done: make(chan struct{})}
	if parameters != nil {
		raw, err := protojson.Marshal(parameters) // the request's prompt payload bytes
		if err != nil {
			return nil, fmt.Errorf("routing: marshal parameters: %w", err)
		}
		p.payload = raw
	}
	s := b.shards[b.nextID.Add(1)%uint64(len(b.shards))] // shard = request id % n_gpus; no exchange
	select {
	case s.queue <- p:
	case <-b.stop:
		return nil, errors.New("routing stage is shut down")
	case <-ctx.Done():
		return nil, ctx.Err()
	}
	select {
	case <-p.done:
	case <-ctx.Done():
		return nil, ctx.Err()
	}
	if p.err != nil { // plain error: server.go:37-40 returns it verbatim (gRPC Unknown)
		return nil, fmt.Errorf("routing failed: %w", p.err)
	}
	return b.inner.ExecuteTool(context.WithValue(ctx, decisionKey{}, p.decision), toolName, parameters, secretId, metadata)
}

func round16(n int) int { return (n + 15) &^ 15 }

func (b *BatchingService) flusher(s *shard) {
	defer b.wg.Done()
	free := make([]int, 0, len(s.arenas))
	for i := range s.arenas {
		free = append(free, i)
	}
	var inflight []int // submitted slots, oldest first
	finishOldest := func() {
		slot := inflight[0]
		inflight = inflight[1:]
		a := &s.arenas[slot]
		err := s.r.Wait(uint32(slot))
		if err != nil {
			b.healthy.Store(false)
		}
		for i, p := range a.reqs {
			p.err = err
			if err == nil {
				p.decision = a.out[i]
			}
			close(p.done)
		}
		a.reqs = a.reqs[:0]
		free = append(free, slot)
	}
	var carry *pending // a request that did not fit the previous arena
	for {
		var first *pending
		if carry != nil {
			first, carry = carry, nil
		} else if len(inflight) > 0 {
			select {
			case first = <-s.queue:
			default: // nothing new: finish the oldest batch now
				finishOldest()
				continue
			}
		} else {
			select {
			case first = <-s.queue:
			case <-b.stop:
				select { // drain what is already queued before leaving
				case first = <-s.queue:
				default:
					return
				}
			}
		}
		if len(free) == 0 {
			finishOldest()
		}
		slot := free[len(free)-1]
		free = free[:len(free)-1]
		a := &s.arenas[slot]
		at, nameAt, n := 0, 0, 0
		add := func(p *pending) bool {
			sz := round16(len(p.payload))
			if at+sz > b.cfg.MaxPayloadBytes || nameAt+len(p.name) > len(a.names) {
				return false
			}
			copy(a.payload[at:], p.payload) // one copy per request into the pinned arena
			a.payOff[n], a.payLen[n] = uint64(at), uint32(len(p.payload))
			a.nameOff[n] = uint32(nameAt)
			copy(a.names[nameAt:], p.name)
			a.reqs = append(a.reqs, p)
			at, nameAt, n = at+sz, nameAt+len(p.name), n+1
			return true
		}
		add(first)
		if !b.cfg.HoldBatches {
			// work-conserving: take what is already queued, do not wait for more
		take:
			for n < b.cfg.MaxBatch {
				select {
				case p := <-s.queue:
					if !add(p) {
						carry = p
						break take
					}
				default:
					break take
				}
			}
		}
		deadline := time.NewTimer(b.cfg.MaxWait)
		idle := time.NewTimer(b.cfg.IdleFlush)
	collect:
		for b.cfg.HoldBatches && n < b.cfg.MaxBatch {
			select {
			case p := <-s.queue:
				if !add(p) {
					carry = p
					break collect
				}
				if !idle.Stop() {
					select {
					case <-idle.C:
					default:
					}
				}
				idle.Reset(b.cfg.IdleFlush)
			case <-idle.C:
				break collect
			case <-deadline.C:
				break collect
			case <-b.stop:
				break collect
			}
		}
		deadline.Stop()
		idle.Stop()
		a.payOff[n], a.nameOff[n] = uint64(at), uint32(nameAt)
		if err := s.r.Submit(uint32(slot), a.payload, a.payOff[:n+1], a.payLen[:n], a.names, a.nameOff[:n+1], a.out[:n]); err != nil {
			b.healthy.Store(false) // fail exactly this batch's requests
			for _, p := range a.reqs {
				p.err = err
				close(p.done)
			}
			a.reqs = a.reqs[:0]
			free = append(free, slot)
			continue
		}
		inflight = append(inflight, slot)
	}
}
