// batch.go — the Go side of the B200 batched-dispatch path.
//
// Drop this file into edegens/ptype's cluster/ directory (package cluster) next to rpc.go and
// registry.go.  It adds a Dispatcher (owner of the GPU context and of the registry snapshot
// image) and a BatchClient whose CallBatch routes and executes N calls exactly as N sequential
// (*Client).Call invocations would (cluster/rpc.go:59-67), through the C ABI declared in
// include/ptype_b200.h.
//
// NOT COMPILED IN THIS REPOSITORY: the build image has no Go toolchain (`go version`: not
// found).  The same ABI is exercised here through Python ctypes (ptype_b200/_abi.py) and
// through the C++ mirror of this file (ptype_b200/host/cluster.hpp, tests/cpp/).
//
// cgo rules this file keeps: records handed to C are plain old data without Go pointers;
// nothing Go-owned is retained by C after a call returns; C strings are freed here.
package cluster

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -L${SRCDIR}/../../ptype_b200 -lptype_b200
#include <stdlib.h>
#include "ptype_b200.h"
*/
import "C"

import (
	"context"
	"errors"
	"fmt"
	"net/rpc"
	"sort"
	"sync"
	"sync/atomic"
	"time"
	"unsafe"
)

// BatchReply mirrors ptype_reply (16 bytes).
type BatchReply struct {
	Value  int64
	Node   uint32 // global node id in the snapshot; Dispatcher.NodeOf resolves it
	Status uint32 // enum ptype_status
}

// Per-call statuses (enum ptype_status).
const (
	StatusOK = iota
	StatusNoService
	StatusNoClient
	StatusIllFormed
	StatusCantFindService
	StatusCantFindMethod
	StatusBadIndex
	StatusRangeLimit
)

// StatusError is what (*Client).Call would have returned for a per-call status.
func StatusError(status uint32, serviceMethod string) error {
	switch status {
	case StatusOK:
		return nil
	case StatusNoService, StatusNoClient:
		return ErrNoClientAvailable // rpc.go:16,62-64
	case StatusIllFormed:
		return rpc.ServerError("rpc: service/method request ill-formed: " + serviceMethod)
	case StatusCantFindService:
		return rpc.ServerError("rpc: can't find service " + serviceMethod)
	case StatusCantFindMethod:
		return rpc.ServerError("rpc: can't find method " + serviceMethod)
	case StatusBadIndex:
		return errors.New("runtime error: index out of range")
	default:
		return fmt.Errorf("ptype_b200: status %d", status)
	}
}

func abiError(where string, rc C.int) error {
	if rc == 0 {
		return nil
	}
	return fmt.Errorf("%s: %s: %s", where, C.GoString(C.ptype_b200_strerror(rc)),
		C.GoString(C.ptype_b200_last_error()))
}

// Dispatcher owns the device context and the whole-registry snapshot shared by every
// BatchClient of this process: per service, the node list its client last accepted.
// Publishing a rebuilt image is the device form of handleNewNodes' swap (rpc.go:233-236).
type Dispatcher struct {
	ctx       *C.ptype_ctx
	localAddr string
	maxConn   int

	mu      sync.Mutex
	lists   map[string][]Node
	exposed map[string][]string
	image   unsafe.Pointer
	salt    uint64
	epoch   uint64
}

// NewDispatcher opens CUDA device `device`.  localAddr and maxConnections are the inputs of
// selectNodes/hashConnectionIndex (rpc.go:252-270): Cluster.localAddr (cluster.go:69,82) and
// ConnConfig.MaxConnections.
func NewDispatcher(localAddr string, maxConnections, device int) (*Dispatcher, error) {
	d := &Dispatcher{localAddr: localAddr, maxConn: maxConnections,
		lists: map[string][]Node{}, exposed: map[string][]string{}}
	dev := C.int(device)
	if err := abiError("ptype_b200_open", C.ptype_b200_open(&dev, 1, &d.ctx)); err != nil {
		return nil, err
	}
	return d, nil
}

func (d *Dispatcher) Close() {
	d.mu.Lock()
	defer d.mu.Unlock()
	if d.image != nil {
		C.ptype_b200_image_free(d.image)
		d.image = nil
	}
	if d.ctx != nil {
		C.ptype_b200_close(d.ctx)
		d.ctx = nil
	}
}

// Expose records that the nodes of `service` ran rpc.Register(new(<receiver>))
// (example/calculator/server/server.go:16-17).
func (d *Dispatcher) Expose(service, receiver string) {
	d.mu.Lock()
	d.exposed[service] = append(d.exposed[service], receiver)
	d.mu.Unlock()
}

// SetNodes installs `nodes` (in WatchService order) as the service's list and publishes the
// next snapshot epoch to the GPU.
func (d *Dispatcher) SetNodes(service string, nodes []Node) error {
	d.mu.Lock()
	defer d.mu.Unlock()
	d.lists[service] = append([]Node(nil), nodes...)
	return d.publishLocked()
}

func (d *Dispatcher) publishLocked() error {
	cAddr := C.CString(d.localAddr)
	defer C.free(unsafe.Pointer(cAddr))
	var b *C.ptype_builder
	if err := abiError("builder_new", C.ptype_b200_builder_new(cAddr, C.int(d.maxConn), &b)); err != nil {
		return err
	}
	defer C.ptype_b200_builder_free(b)
	services := make([]string, 0, len(d.lists))
	for s := range d.lists {
		services = append(services, s)
	}
	sort.Strings(services)
	for _, s := range services {
		cs := C.CString(s)
		C.ptype_b200_builder_declare(b, cs)
		for i, n := range d.lists[s] {
			// synthetic node names keep the watch's order under the etcd byte sort
			cn, ca := C.CString(fmt.Sprintf("%08d", i)), C.CString(n.Address)
			rc := C.ptype_b200_builder_register(b, cs, cn, ca, C.int(n.Port))
			C.free(unsafe.Pointer(cn))
			C.free(unsafe.Pointer(ca))
			if err := abiError("builder_register", rc); err != nil {
				C.free(unsafe.Pointer(cs))
				return err
			}
		}
		for _, r := range d.exposed[s] {
			cr := C.CString(r)
			C.ptype_b200_builder_expose(b, cs, cr)
			C.free(unsafe.Pointer(cr))
		}
		C.free(unsafe.Pointer(cs))
	}
	var blob unsafe.Pointer
	var size C.size_t
	if err := abiError("builder_finish", C.ptype_b200_builder_finish(b, C.uint64_t(d.epoch+1), &blob, &size)); err != nil {
		return err
	}
	if err := abiError("publish_image", C.ptype_b200_publish_image(d.ctx, blob, size)); err != nil {
		C.ptype_b200_image_free(blob)
		return err
	}
	if d.image != nil {
		C.ptype_b200_image_free(d.image)
	}
	d.image = blob
	d.epoch++
	d.salt = uint64((*C.ptype_image_header)(blob).key_salt)
	return nil
}

// UpdateService applies one watch event in place: a node joined (+1) or left (-1) `service`.
// The table is patched on the GPU (ptype_b200_update_service); nothing is rebuilt or uploaded.
// The Dispatcher's own node lists must be kept in step by the caller (SetNodes does both).
func (d *Dispatcher) UpdateService(service string, delta int) error {
	d.mu.Lock()
	defer d.mu.Unlock()
	cs := C.CString(service)
	defer C.free(unsafe.Pointer(cs))
	if err := abiError("ptype_b200_update_service", C.ptype_b200_update_service(d.ctx, cs, C.int(delta))); err != nil {
		return err
	}
	d.epoch++
	return nil
}

// AssignSeq fills the seq field of n device-resident request records the way n sequential
// Client.Call()s, each on the Client of its record's service, would have drawn them: every
// service's counter (connectionBalancer.seq, rpc.go:129) lives on the GPU, advances only for
// calls that find a non-empty client list (rpc.go:178-181), and follows its service through
// SetNodes / UpdateService.  dReqs is device memory from ptype_b200_device_alloc.
func (d *Dispatcher) AssignSeq(dReqs unsafe.Pointer, n int) error {
	return abiError("ptype_b200_assign_seq",
		C.ptype_b200_assign_seq(d.ctx, 0, (*C.ptype_req)(dReqs), C.size_t(n), 0))
}

// MixedCall mirrors ptype_req (48 bytes): one call of a batch that mixes services.  Seq is
// filled in by CallMixed.
type MixedCall struct {
	SvcKey, Seq  uint64
	Method, Recv uint32
	A0, A1, A2   int64
}

// CallMixed issues calls[0], calls[1], ... as the process would through one Client per service,
// in that order: sequence numbers are drawn on the GPU from each service's own counter
// (AssignSeq), then the batch is dispatched.  On return calls[j].Seq holds the number drawn
// (0 for a call that found no client) and replies[j] the outcome.
func (d *Dispatcher) CallMixed(calls []MixedCall, replies []BatchReply) error {
	if len(calls) == 0 {
		return nil
	}
	if len(replies) < len(calls) {
		return errors.New("ptype_b200: replies shorter than calls")
	}
	n := C.size_t(len(calls))
	var dReq, dRep unsafe.Pointer
	if err := abiError("ptype_b200_device_alloc", C.ptype_b200_device_alloc(d.ctx, 0, n*48, &dReq)); err != nil {
		return err
	}
	defer C.ptype_b200_device_free(d.ctx, 0, dReq)
	if err := abiError("ptype_b200_device_alloc", C.ptype_b200_device_alloc(d.ctx, 0, n*16, &dRep)); err != nil {
		return err
	}
	defer C.ptype_b200_device_free(d.ctx, 0, dRep)
	if err := abiError("ptype_b200_memcpy_h2d",
		C.ptype_b200_memcpy_h2d(d.ctx, 0, dReq, unsafe.Pointer(&calls[0]), n*48)); err != nil {
		return err
	}
	if err := d.AssignSeq(dReq, len(calls)); err != nil {
		return err
	}
	if err := abiError("ptype_b200_dispatch", C.ptype_b200_dispatch(d.ctx, 0, (*C.ptype_req)(dReq), n,
		(*C.ptype_reply)(dRep), C.PTYPE_F_DEVICE_PTRS)); err != nil {
		return err
	}
	if err := abiError("ptype_b200_memcpy_d2h",
		C.ptype_b200_memcpy_d2h(d.ctx, 0, unsafe.Pointer(&replies[0]), dRep, n*16)); err != nil {
		return err
	}
	return abiError("ptype_b200_memcpy_d2h",
		C.ptype_b200_memcpy_d2h(d.ctx, 0, unsafe.Pointer(&calls[0]), dReq, n*48))
}

// Seq reads a service's counter: the last value drawn.
func (d *Dispatcher) Seq(service string) (uint64, error) {
	cs := C.CString(service)
	defer C.free(unsafe.Pointer(cs))
	var v C.uint64_t
	err := abiError("ptype_b200_seq_get", C.ptype_b200_seq_get(d.ctx, 0, cs, &v))
	return uint64(v), err
}

// SelectedNodes is getSelectedNodes() (rpc.go:293-297) read back from the snapshot.
func (d *Dispatcher) SelectedNodes(service string) []Node {
	d.mu.Lock()
	defer d.mu.Unlock()
	if d.image == nil {
		return nil
	}
	cs := C.CString(service)
	defer C.free(unsafe.Pointer(cs))
	ids := make([]C.uint32_t, 4096)
	var k C.uint32_t
	if C.ptype_b200_image_selected(d.image, cs, &ids[0], C.size_t(len(ids)), &k) != 0 {
		return nil
	}
	if int(k) > len(ids) {
		ids = make([]C.uint32_t, int(k))
		C.ptype_b200_image_selected(d.image, cs, &ids[0], C.size_t(len(ids)), &k)
	}
	out := make([]Node, 0, int(k))
	for i := 0; i < int(k); i++ {
		out = append(out, d.nodeOfLocked(uint32(ids[i])))
	}
	return out
}

// NodeOf resolves a BatchReply.Node.
func (d *Dispatcher) NodeOf(id uint32) Node {
	d.mu.Lock()
	defer d.mu.Unlock()
	return d.nodeOfLocked(id)
}

func (d *Dispatcher) nodeOfLocked(id uint32) Node {
	var addr [256]C.char
	var port C.int
	if d.image == nil || C.ptype_b200_image_node(d.image, C.uint32_t(id), &addr[0], 256, &port, nil, 0, nil, 0) != 0 {
		return Node{}
	}
	return Node{Address: C.GoString(&addr[0]), Port: int(port)}
}

// BatchClient is a Client whose calls run on the GPU.  It keeps the reference's balancer
// sequence (one atomic counter per client, rpc.go:129,181) and its update rule (latest
// non-empty list after DebounceTime of quiet, rpc.go:197-224).
type BatchClient struct {
	gpu     *Dispatcher
	service string
	cfg     *ConnConfig
	seq     uint64

	cancel    func()
	errChan   chan error
	updated   chan struct{}
	waitGroup sync.WaitGroup
}

// NewBatchClient mirrors newClient/newConnectionBalancer (rpc.go:45-57,141-170).
func NewBatchClient(gpu *Dispatcher, serviceName string, r Registry, cfg *ConnConfig) (*BatchClient, error) {
	if cfg == nil {
		cfg = DefaultConnConfig
	}
	if cfg.MaxConnections != gpu.maxConn {
		return nil, errors.New("ptype_b200: ConnConfig.MaxConnections must equal the Dispatcher's")
	}
	ctx, cancel := context.WithCancel(context.Background())
	c := &BatchClient{gpu: gpu, service: serviceName, cfg: cfg, cancel: cancel,
		errChan: make(chan error, 1), updated: make(chan struct{}, 5)}
	nodesChan := r.WatchService(ctx, serviceName)
	var initial []Node
	select {
	case initial = <-nodesChan:
	case <-time.After(cfg.InitialNodeTimeout):
		cancel()
		return nil, fmt.Errorf("no initial nodes provided for %v", serviceName)
	}
	if err := gpu.SetNodes(serviceName, initial); err != nil {
		cancel()
		return nil, err
	}
	c.waitGroup.Add(1)
	go c.watchForNewNodes(ctx, nodesChan)
	return c, nil
}

func (c *BatchClient) watchForNewNodes(ctx context.Context, nodesChan chan []Node) {
	defer c.waitGroup.Done()
	var newNodes []Node
	for {
		select {
		case nodes := <-nodesChan:
			if len(nodes) == 0 {
				continue
			}
			newNodes = nodes
		case <-time.After(c.cfg.DebounceTime):
			if newNodes == nil {
				continue
			}
			if err := c.gpu.SetNodes(c.service, newNodes); err != nil {
				select {
				case c.errChan <- err:
				default:
				}
			} else {
				select {
				case c.updated <- struct{}{}:
				default:
				}
			}
			newNodes = nil
		case <-ctx.Done():
			return
		}
	}
}

func (c *BatchClient) ConnectionErrs() chan error { return c.errChan }

func (c *BatchClient) Close() error {
	c.cancel()
	c.waitGroup.Wait()
	close(c.errChan)
	return nil
}

// Call is (*Client).Call for one calculator- or prime-style call.
func (c *BatchClient) Call(serviceMethod string, args [3]int64, reply *int64) error {
	rep := make([]BatchReply, 1)
	if err := c.CallBatch(serviceMethod, [][3]int64{args}, rep); err != nil {
		return err
	}
	if err := StatusError(rep[0].Status, serviceMethod); err != nil {
		return err // *reply untouched, like net/rpc on error
	}
	*reply = rep[0].Value
	return nil
}

// CallBatch issues len(args) calls of one method.  The sequence range is reserved with ONE
// atomic add, so routing equals len(args) sequential Call()s.  Failed calls are re-issued with
// fresh sequence numbers up to cfg.Retries more times (withRetry, rpc.go:107-116, bounded).
func (c *BatchClient) CallBatch(serviceMethod string, args [][3]int64, replies []BatchReply) error {
	n := len(args)
	if n == 0 {
		return nil
	}
	if len(replies) < n {
		return errors.New("ptype_b200: replies shorter than args")
	}
	csm := C.CString(serviceMethod)
	var method, recv C.uint32_t
	C.ptype_b200_method_key(csm, &method, &recv)
	C.free(unsafe.Pointer(csm))
	c.gpu.mu.Lock()
	salt := c.gpu.salt
	c.gpu.mu.Unlock()
	cservice := C.CString(c.service)
	key := C.ptype_b200_service_key(cservice, C.uint64_t(salt))
	C.free(unsafe.Pointer(cservice))

	// C-owned pinned staging: Go memory is never handed to the asynchronous copy engine.
	reqs := (*[1 << 30]C.ptype_req)(C.ptype_b200_alloc_pinned(C.size_t(n) * C.sizeof_ptype_req))[:n:n]
	out := (*[1 << 30]C.ptype_reply)(C.ptype_b200_alloc_pinned(C.size_t(n) * C.sizeof_ptype_reply))[:n:n]
	failed := (*[1 << 30]C.uint32_t)(C.ptype_b200_alloc_pinned(C.size_t(n) * 4))[:n:n]
	defer C.ptype_b200_free_pinned(unsafe.Pointer(&reqs[0]))
	defer C.ptype_b200_free_pinned(unsafe.Pointer(&out[0]))
	defer C.ptype_b200_free_pinned(unsafe.Pointer(&failed[0]))

	todo := make([]int, n)
	for j := range todo {
		todo[j] = j
	}
	for attempt := 0; ; attempt++ {
		m := uint64(len(todo))
		base := atomic.AddUint64(&c.seq, m) - m // rpc.go:181, batched: first call sees base+1
		for t, j := range todo {
			reqs[t] = C.ptype_req{svc_key: key, seq: C.uint64_t(base + 1 + uint64(t)), method: method, recv: recv,
				a0: C.int64_t(args[j][0]), a1: C.int64_t(args[j][1]), a2: C.int64_t(args[j][2])}
		}
		var stats C.ptype_batch_stats
		rc := C.ptype_b200_dispatch_ex(c.gpu.ctx, 0, &reqs[0], C.size_t(m), &out[0], &failed[0], C.size_t(m), &stats, 0)
		if err := abiError("ptype_b200_dispatch_ex", rc); err != nil {
			return err
		}
		for t, j := range todo {
			replies[j] = BatchReply{Value: int64(out[t].value), Node: uint32(out[t].node), Status: uint32(out[t].status)}
		}
		if stats.failed_count == 0 || attempt >= c.cfg.Retries {
			return nil
		}
		next := make([]int, 0, int(stats.failed_count))
		for f := 0; f < int(stats.failed_count); f++ {
			next = append(next, todo[int(failed[f])])
		}
		sort.Ints(next)
		todo = next
	}
}

// Coalescer is the dispatch loop in front of the GPU path: goroutines call Call concurrently,
// as they do with (*Client).Call today; one goroutine gathers what arrived within MaxWait (or
// MaxBatch calls), submits one CallBatch per method name and hands every caller its reply.
type Coalescer struct {
	c        *BatchClient
	MaxBatch int
	MaxWait  time.Duration
	in       chan *pendingCall
	done     chan struct{}
}

type pendingCall struct {
	method string
	args   [3]int64
	rep    BatchReply
	err    error
	ready  chan struct{}
}

func NewCoalescer(c *BatchClient) *Coalescer {
	co := &Coalescer{c: c, MaxBatch: 65536, MaxWait: 200 * time.Microsecond,
		in: make(chan *pendingCall, 1<<16), done: make(chan struct{})}
	go co.run()
	return co
}

func (co *Coalescer) Close() { close(co.in); <-co.done }

// Call blocks until the call has been dispatched; results and errors are those of BatchClient.Call.
func (co *Coalescer) Call(serviceMethod string, args [3]int64, reply *int64) error {
	p := &pendingCall{method: serviceMethod, args: args, ready: make(chan struct{})}
	co.in <- p
	<-p.ready
	if p.err != nil {
		return p.err
	}
	if err := StatusError(p.rep.Status, serviceMethod); err != nil {
		return err
	}
	*reply = p.rep.Value
	return nil
}

func (co *Coalescer) run() {
	defer close(co.done)
	for first := range co.in {
		groups := map[string][]*pendingCall{first.method: {first}}
		n := 1
		timer := time.NewTimer(co.MaxWait)
	gather:
		for n < co.MaxBatch {
			select {
			case p, ok := <-co.in:
				if !ok {
					break gather
				}
				groups[p.method] = append(groups[p.method], p)
				n++
			case <-timer.C:
				break gather
			}
		}
		timer.Stop()
		for method, ps := range groups {
			args := make([][3]int64, len(ps))
			for i, p := range ps {
				args[i] = p.args
			}
			reps := make([]BatchReply, len(ps))
			err := co.c.CallBatch(method, args, reps)
			for i, p := range ps {
				p.rep, p.err = reps[i], err
				close(p.ready)
			}
		}
	}
}
