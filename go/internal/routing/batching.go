package routing

// UNVERIFIED SOURCE (no Go toolchain where this was written) — a transliteration of
// polykey_service_b200/host/batching_service.cpp, which is compiled, tested (CPU stand-in and
// CUDA back-end, ThreadSanitizer) and measured.  Keep the two in step.
//
// Shape (see batching_service.hpp for the reasoning and the numbers):
//   - every ExecuteTool call packs ITS OWN request: it takes the next record index of the open
//     batch with one atomic add, copies its payload into that index's fixed slot of the pinned
//     arena and its tool name into the index's 64-byte record, and commits by storing its length;
//   - ONE flusher goroutine per GPU seals the open batch, opens the next arena so producers
//     carry on, waits for the last commits while it spreads the records into the arrays of
//     include/pk_route.h, crosses cgo once per batch (pk_submit / pk_wait), and publishes the
//     whole batch by closing one channel; callers read their own 24-byte decision;
//   - a request that does not fit (payload over the arena, name over MaxNameBytes) never reaches
//     the device: the first is an error for that request alone, the second is answered by the
//     inner service without a decision (it is the switch's default branch whatever its bytes,
//     mock.go:60-64);
//   - only a device failure (PK_ERR_CUDA / PK_ERR_NO_DEVICE) turns Healthy() false; a batch the
//     device refuses fails its own requests.

import (
	"context"
	"errors"
	"runtime"
	"sync"
	"sync/atomic"
	"time"

	"github.com/spounge-ai/polykey-service/internal/service"
	cmn "github.com/spounge-ai/spounge-proto/gen/go/common/v2"
	pk "github.com/spounge-ai/spounge-proto/gen/go/polykey/v2"
	"google.golang.org/protobuf/encoding/protojson"
	"google.golang.org/protobuf/types/known/structpb"
)

// Config: the knobs follow the reference's pattern of env vars read in main
// (cmd/polykey/main.go:57-60): PK_MAX_BATCH, PK_MAX_WAIT_US, PK_SLOT_BYTES, PK_DEVICES.
type Config struct {
	MaxBatch     int           // requests per flush (<= 65536)
	SlotBytes    int           // payload slot of one request (multiple of 16); a larger payload takes several
	MaxNameBytes int           // longest tool name that is routed (default 256)
	MaxWait      time.Duration // hold-the-batch policy: deadline from the first request of a batch
	HoldBatches  bool          // false (default): work-conserving — a free slot takes what is there now
	Devices      []int         // one flusher, one Router, one table replica per entry; listing a GPU twice gives it two
	                           // shards — two open batches, hence two reservation counters: measured with the C++
	                           // mirror, one GPU takes 1.5e7 requests/s of 2 KB with one shard and 2.05e7 with two
	Slots        int           // batches in flight on a GPU
}

const (
	sealed      = uint64(1) << 63
	uncommitted = ^uint32(0)
	placeholder = ^uint32(0)
	unset       = ^uint32(0)
	inlineName  = 56
)

// rec is what a producer says about its request: one cache line of its own.
type rec struct {
	length  atomic.Uint32 // payload bytes; uncommitted until copied, and again once picked up
	nameLen uint32        // placeholder: an index covered by the bigger payload before it
	name    [inlineName]byte
}

type arena struct {
	payload   []byte     // pinned (Router.HostAlloc): request i at payload[i*SlotBytes:]
	out       []Decision // pinned
	longNames []byte     // names over inlineName bytes, MaxNameBytes apart
	recs      []rec
	// the arrays pk_submit takes; filled by the flusher (pk_submit consumes them before it returns)
	payOff  []uint64
	payLen  []uint32
	nameOff []uint32
	names   []byte
	state   atomic.Uint64 // [sealed:1][slots taken:63]
	finalN  atomic.Uint32 // slots taken successfully, published by whoever closed the batch
	done    chan struct{} // closed when `out` (or err) is ready; a fresh one per generation
	err     error
	n       int // flusher-private
	picked  int
}

type shard struct {
	r       *Router
	arenas  []*arena
	open    atomic.Pointer[arena] // nil: every arena is busy
	changed atomic.Pointer[chan struct{}]
	closed  atomic.Bool
	kick    chan struct{} // wakes an idle flusher
}

// BatchingService implements service.Service (internal/service/service.go:13-15).  It is
// handed to server.NewServer where NewMockService() is today (cmd/polykey/main.go:85-86).
type BatchingService struct {
	inner    service.Service
	cfg      Config
	shards   []*shard
	nextID   atomic.Uint64
	healthy  atomic.Bool
	stopping atomic.Bool
	wg       sync.WaitGroup
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
	if cfg.MaxBatch <= 0 || cfg.MaxBatch > 65536 || len(cfg.Devices) == 0 {
		return nil, errors.New("routing: bad config")
	}
	if cfg.Slots <= 0 {
		cfg.Slots = 3
	}
	if cfg.MaxNameBytes <= 0 {
		cfg.MaxNameBytes = 256
	}
	cfg.SlotBytes = (cfg.SlotBytes + 15) &^ 15
	if cfg.SlotBytes == 0 {
		cfg.SlotBytes = 8192
	}
	b := &BatchingService{inner: inner, cfg: cfg}
	b.healthy.Store(true)
	for _, dev := range cfg.Devices {
		r, err := NewRouter(dev, uint32(cfg.MaxBatch), uint64(cfg.MaxBatch*cfg.SlotBytes), uint32(cfg.Slots))
		if err != nil {
			b.Shutdown()
			return nil, err
		}
		s := &shard{r: r, kick: make(chan struct{}, 1)}
		ch := make(chan struct{})
		s.changed.Store(&ch)
		// more arenas than device slots: an arena stays taken until every caller has picked its
		// record up, and that must not hold a device slot hostage
		for k := 0; k < 2*cfg.Slots+1; k++ {
			a := &arena{
				payload:   r.HostAlloc(cfg.MaxBatch*cfg.SlotBytes + 64),
				out:       r.HostAllocDecisions(cfg.MaxBatch),
				longNames: make([]byte, cfg.MaxBatch*cfg.MaxNameBytes),
				recs:      make([]rec, cfg.MaxBatch),
				payOff:    make([]uint64, cfg.MaxBatch+1),
				payLen:    make([]uint32, cfg.MaxBatch),
				nameOff:   make([]uint32, cfg.MaxBatch+1),
				names:     make([]byte, cfg.MaxBatch*cfg.MaxNameBytes),
			}
			if a.payload == nil || a.out == nil {
				r.Close() // not yet in b.shards: Shutdown below would not reach it
				b.Shutdown()
				return nil, errors.New("routing: pinned allocation failed")
			}
			for i := range a.recs {
				a.recs[i].length.Store(uncommitted)
			}
			a.state.Store(sealed)
			s.arenas = append(s.arenas, a)
		}
		b.shards = append(b.shards, s)
	}
	for _, s := range b.shards {
		b.wg.Add(1)
		go b.flusher(s)
	}
	return b, nil
}

func (b *BatchingService) Healthy() bool { return b.healthy.Load() }

// Shutdown drains and stops the flushers; call it before srv.GracefulStop()
// (cmd/polykey/main.go:118-119).  Calls that arrive afterwards fail at once.
func (b *BatchingService) Shutdown() {
	if b.stopping.Swap(true) {
		return
	}
	for _, s := range b.shards {
		s.wake()
	}
	b.wg.Wait()
	for _, s := range b.shards {
		s.r.Close()
	}
}

func (s *shard) wake() {
	select {
	case s.kick <- struct{}{}:
	default:
	}
}

// announce tells producers waiting for an arena that `open` has changed.
func (s *shard) announce() {
	ch := make(chan struct{})
	old := s.changed.Swap(&ch)
	close(*old)
}

type ticket struct {
	a     *arena
	idx   int
	slots int
}

// reserve takes the next record index of the open batch (k of them when the payload spans k
// slots) with ONE atomic add, copies the request in and commits.  An add cannot be taken back,
// so a request that does not fit has still moved the counter; that is harmless because nothing
// fits after it either, and because the number of slots that WERE taken is published once, by
// exactly one party: the producer whose request was the first not to fit, or, if the batch never
// overflowed, whoever set the sealed bit first.
func (b *BatchingService) reserve(ctx context.Context, s *shard, name string, payload []byte) (ticket, error) {
	slot := uint64(b.cfg.SlotBytes)
	k := uint64(1)
	if uint64(len(payload)) > slot {
		k = (uint64(len(payload)) + slot - 1) / slot
	}
	capacity := uint64(b.cfg.MaxBatch)
	for {
		changed := *s.changed.Load()
		if a := s.open.Load(); a != nil {
			seen := a.state.Load()
			if seen&sealed == 0 && seen+k <= capacity { // do not touch a batch that cannot take us
				old := a.state.Add(k) - k
				if old&sealed == 0 {
					if old+k <= capacity {
						i := int(old)
						copy(a.payload[i*b.cfg.SlotBytes:], payload)
						r := &a.recs[i]
						r.nameLen = uint32(len(name))
						if len(name) <= inlineName {
							copy(r.name[:], name)
						} else {
							copy(a.longNames[i*b.cfg.MaxNameBytes:], name)
						}
						for j := 1; j < int(k); j++ { // indices our payload runs over: empty requests
							a.recs[i+j].nameLen = placeholder
							a.recs[i+j].length.Store(0)
						}
						r.length.Store(uint32(len(payload))) // commit
						if i == 0 {
							s.wake() // a batch has been started
						}
						return ticket{a, i, int(k)}, nil
					}
					if old <= capacity { // the first request that did not fit: `old` slots were taken
						a.finalN.Store(uint32(old))
						a.state.Or(sealed)
						s.wake()
					}
				}
			} else if seen&sealed == 0 && seen <= capacity {
				// no room for this request in what is left of the batch: close it for the flusher
				if old := a.state.Or(sealed); old&sealed == 0 && old <= capacity {
					a.finalN.Store(uint32(old))
				}
				s.wake()
			}
		}
		if s.closed.Load() {
			return ticket{}, errors.New("routing stage is shut down")
		}
		select { // every arena is busy, or the open one has just been sealed: wait for the next
		case <-changed:
		case <-ctx.Done():
			return ticket{}, ctx.Err()
		}
	}
}

// ExecuteTool — same signature as service.Service (internal/service/service.go:13-15).
func (b *BatchingService) ExecuteTool(ctx context.Context, toolName string, parameters *structpb.Struct,
	secretId *string, metadata *cmn.Metadata) (*pk.ExecuteToolResponse, error) {
	if b.stopping.Load() || len(b.shards) == 0 {
		return nil, errors.New("routing stage is shut down")
	}
	if !b.healthy.Load() {
		return nil, errors.New("routing stage is unhealthy")
	}
	// A name this long equals no table entry: it is the switch's default branch (mock.go:60-64)
	// whatever the routing stage would say, so answer it without one.
	if len(toolName) > b.cfg.MaxNameBytes {
		return b.inner.ExecuteTool(ctx, toolName, parameters, secretId, metadata)
	}
	var payload []byte
	if parameters != nil {
		var err error
		if payload, err = protojson.Marshal(parameters); err != nil { // the request's prompt payload bytes
			return nil, err
		}
	}
	if len(payload) > b.cfg.MaxBatch*b.cfg.SlotBytes {
		return nil, errors.New("request payload exceeds the routing stage's arena")
	}
	id := b.nextID.Add(1) - 1
	s := b.shards[id%uint64(len(b.shards))] // shard by request id, no exchange
	t, err := b.reserve(ctx, s, toolName, payload)
	if err != nil {
		return nil, err
	}
	<-t.a.done // the batch is published by closing this channel; never skipped: the arena waits for our pick-up
	batchErr := t.a.err
	var d Decision
	if batchErr == nil {
		d = t.a.out[t.idx]
	}
	for j := t.slots - 1; j >= 0; j-- { // picked up (placeholders too): our last access to the arena
		t.a.recs[t.idx+j].length.Store(uncommitted)
	}
	if batchErr != nil {
		return nil, batchErr // plain error: server.go:37-40 returns it verbatim (gRPC Unknown)
	}
	return b.inner.ExecuteTool(context.WithValue(ctx, decisionKey{}, d), toolName, parameters, secretId, metadata)
}

// seal closes batch a to further reservations and returns how many slots were taken (see reserve).
func (b *BatchingService) seal(a *arena) int {
	if old := a.state.Or(sealed); old&sealed == 0 && old <= uint64(b.cfg.MaxBatch) {
		a.finalN.Store(uint32(old)) // we closed it, and it had not overflowed
		return int(old)
	}
	for { // closed (or overflowed) by a producer, who says how many — a few instructions away
		if n := a.finalN.Load(); n != unset {
			return int(n)
		}
		runtime.Gosched()
	}
}

func (s *shard) openArena(a *arena) {
	a.n, a.picked, a.err = 0, 0, nil
	a.done = make(chan struct{})
	a.finalN.Store(unset)
	a.state.Store(0) // publishes done and finalN
	s.open.Store(a)
	s.announce()
}

type flight struct {
	a    *arena
	slot uint32
}

// flusher: one per GPU.  Never touches payload bytes.
func (b *BatchingService) flusher(s *shard) {
	defer b.wg.Done()
	runtime.LockOSThread() // the CUDA context and the polling belong to one OS thread
	var inflight []flight
	var draining []*arena
	free := append([]*arena(nil), s.arenas...)
	freeSlots := make([]uint32, 0, b.cfg.Slots)
	for i := b.cfg.Slots; i > 0; i-- {
		freeSlots = append(freeSlots, uint32(i-1))
	}
	reap := func() {
		keep := draining[:0]
		for _, d := range draining {
			for d.picked < d.n && d.recs[d.picked].length.Load() == uncommitted {
				d.picked++
			}
			if d.picked == d.n {
				free = append(free, d)
			} else {
				keep = append(keep, d)
			}
		}
		draining = keep
	}
	complete := func(a *arena, err error) {
		if err != nil && IsDeviceFailure(err) {
			b.healthy.Store(false) // -> health NOT_SERVING (cmd/polykey/main.go:93-94)
		}
		a.err = err
		close(a.done) // publishes the whole batch
		draining = append(draining, a)
	}
	finishOldest := func() {
		f := inflight[0]
		inflight = inflight[1:]
		complete(f.a, s.r.Wait(f.slot))
		freeSlots = append(freeSlots, f.slot)
	}
	takeFree := func() *arena {
		for {
			reap()
			if n := len(free); n > 0 {
				a := free[n-1]
				free = free[:n-1]
				return a
			}
			if len(inflight) > 0 {
				finishOldest()
				continue
			}
			runtime.Gosched() // all completed, some records not yet picked up
		}
	}

	open := takeFree()
	s.openArena(open)
	var firstSeen time.Time
	for {
		st := open.state.Load()
		if st&^sealed == 0 && st&sealed == 0 {
			firstSeen = time.Time{}
			if len(inflight) > 0 {
				finishOldest() // nothing new: finish the oldest batch now
				continue
			}
			reap()
			if b.stopping.Load() {
				if b.seal(open) == 0 { // shut the door; a producer that slipped in still gets its batch
					break
				}
			} else {
				select { // idle: sleep until a producer starts a batch (or shutdown)
				case <-s.kick:
				case <-time.After(10 * time.Millisecond):
				}
				continue
			}
		}
		if firstSeen.IsZero() {
			firstSeen = time.Now()
		}
		if b.cfg.HoldBatches && !b.stopping.Load() { // hold the batch for more, up to MaxWait from its first request
			for st = open.state.Load(); st&sealed == 0 && st < uint64(b.cfg.MaxBatch) && time.Since(firstSeen) < b.cfg.MaxWait; st = open.state.Load() {
				runtime.Gosched()
			}
		}
		next := takeFree() // before sealing: producers never find the door shut for long
		n := b.seal(open)
		s.openArena(next)
		firstSeen = time.Time{}
		a := open
		open = next
		if n == 0 {
			free = append(free, a)
			continue
		}
		// wait for the last commits, spreading the records into the arrays the C ABI takes
		var namesAt uint32
		var prevEnd uint64
		for i := 0; i < n; i++ {
			r := &a.recs[i]
			l := r.length.Load()
			for l == uncommitted {
				runtime.Gosched()
				l = r.length.Load()
			}
			off := uint64(i * b.cfg.SlotBytes)
			if off < prevEnd {
				off = prevEnd // an index a bigger payload runs over starts where that payload ends
			}
			a.payOff[i], a.payLen[i], a.nameOff[i] = off, l, namesAt
			prevEnd = off + (uint64(l)+15)&^15
			if r.nameLen == placeholder || r.nameLen == 0 {
				continue
			}
			if r.nameLen <= inlineName {
				copy(a.names[namesAt:], r.name[:r.nameLen])
			} else {
				copy(a.names[namesAt:], a.longNames[i*b.cfg.MaxNameBytes:][:r.nameLen])
			}
			namesAt += r.nameLen
		}
		a.payOff[n] = uint64(n * b.cfg.SlotBytes)
		if a.payOff[n] < prevEnd {
			a.payOff[n] = prevEnd
		}
		a.nameOff[n] = namesAt
		a.n = n
		if len(freeSlots) == 0 {
			finishOldest() // a device slot for this batch
		}
		slot := freeSlots[len(freeSlots)-1]
		freeSlots = freeSlots[:len(freeSlots)-1]
		if err := s.r.Submit(slot, a.payload, a.payOff[:n+1], a.payLen[:n], a.names, a.nameOff[:n+1], a.out[:n]); err != nil {
			complete(a, err) // fail exactly this batch's requests
			freeSlots = append(freeSlots, slot)
		} else {
			inflight = append(inflight, flight{a, slot})
		}
	}
	// stopping, the open batch sealed empty: refuse further reservations, finish what is in flight
	s.closed.Store(true)
	s.open.Store(nil)
	s.announce()
	for len(inflight) > 0 {
		finishOldest()
	}
	for deadline := time.Now().Add(5 * time.Second); len(draining) > 0 && time.Now().Before(deadline); reap() {
		runtime.Gosched()
	}
}
