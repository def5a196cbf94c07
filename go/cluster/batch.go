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

// Snapshot publication through the KV store (SURVEY.md 8f N3).  One host builds the registry image
// (SetNodes) and publishes it under store/registry_image/<epoch>; every other host of the
// deployment installs the newest epoch, so all GPUs answer from the same bytes.  Only the
// unchanged KVStore.Put / KVStore.Get (store.go:38-62) are used; etcd values are byte strings,
// so the blob travels as it is.
const imageKeyPrefix = "registry_image/"

// PublishImage stores the image this Dispatcher last built.
func (d *Dispatcher) PublishImage(ctx context.Context, store *KVStore) error {
	d.mu.Lock()
	defer d.mu.Unlock()
	if d.image == nil {
		return errors.New("ptype_b200: no image built yet")
	}
	h := (*C.ptype_image_header)(d.image)
	blob := C.GoBytes(d.image, C.int(h.total_bytes))
	return store.Put(ctx, fmt.Sprintf("%s%020d", imageKeyPrefix, uint64(h.epoch)), string(blob))
}

// LoadLatestImage installs the newest published image on this Dispatcher's GPU and returns its
// epoch.  ErrNoKey when nothing has been published.  A Dispatcher that follows a publisher this
// way does not call SetNodes itself.
func (d *Dispatcher) LoadLatestImage(ctx context.Context, store *KVStore) (uint64, error) {
	vals, err := store.Get(ctx, imageKeyPrefix, WithPrefix(), WithSort(SortByKey, SortDescend), WithLimit(1))
	if err != nil {
		return 0, err
	}
	blob := C.CBytes([]byte(vals[0])) // malloc'ed copy: ptype_b200_image_free is free()
	size := C.size_t(len(vals[0]))
	if err := abiError("ptype_b200_image_validate", C.ptype_b200_image_validate(blob, size)); err != nil {
		C.free(blob)
		return 0, err
	}
	d.mu.Lock()
	defer d.mu.Unlock()
	if err := abiError("ptype_b200_publish_image", C.ptype_b200_publish_image(d.ctx, blob, size)); err != nil {
		C.free(blob)
		return 0, err
	}
	if d.image != nil {
		C.ptype_b200_image_free(d.image)
	}
	d.image = blob
	h := (*C.ptype_image_header)(blob)
	d.epoch, d.salt = uint64(h.epoch), uint64(h.key_salt)
	return d.epoch, nil
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
	var dReq, dRep, dSeq unsafe.Pointer
	if err := abiError("ptype_b200_device_alloc", C.ptype_b200_device_alloc(d.ctx, 0, n*48, &dReq)); err != nil {
		return err
	}
	defer C.ptype_b200_device_free(d.ctx, 0, dReq)
	if err := abiError("ptype_b200_device_alloc", C.ptype_b200_device_alloc(d.ctx, 0, n*16, &dRep)); err != nil {
		return err
	}
	defer C.ptype_b200_device_free(d.ctx, 0, dRep)
	if err := abiError("ptype_b200_device_alloc", C.ptype_b200_device_alloc(d.ctx, 0, n*8, &dSeq)); err != nil {
		return err
	}
	defer C.ptype_b200_device_free(d.ctx, 0, dSeq)
	if err := abiError("ptype_b200_memcpy_h2d",
		C.ptype_b200_memcpy_h2d(d.ctx, 0, dReq, unsafe.Pointer(&calls[0]), n*48)); err != nil {
		return err
	}
	// The numbers go to an array beside the records and the dispatch reads them there: storing
	// them into the 48-byte records would cost the GPU a read and a write of a sector per call.
	if err := abiError("ptype_b200_assign_seq_array", C.ptype_b200_assign_seq_array(d.ctx, 0,
		(*C.ptype_req)(dReq), n, (*C.uint64_t)(dSeq), C.PTYPE_F_ASYNC)); err != nil {
		return err
	}
	if err := abiError("ptype_b200_dispatch_seq", C.ptype_b200_dispatch_seq(d.ctx, 0, (*C.ptype_req)(dReq),
		(*C.uint64_t)(dSeq), n, (*C.ptype_reply)(dRep), 0)); err != nil {
		return err
	}
	if err := abiError("ptype_b200_memcpy_d2h",
		C.ptype_b200_memcpy_d2h(d.ctx, 0, unsafe.Pointer(&replies[0]), dRep, n*16)); err != nil {
		return err
	}
	seq := make([]uint64, len(calls))
	if err := abiError("ptype_b200_memcpy_d2h",
		C.ptype_b200_memcpy_d2h(d.ctx, 0, unsafe.Pointer(&seq[0]), dSeq, n*8)); err != nil {
		return err
	}
	for j := range calls {
		calls[j].Seq = seq[j]
	}
	return nil
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

	poolMu sync.Mutex
	pool   []*Batch

	coalescerOnce sync.Once
	coalescer     *Coalescer // started by the first Go()
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
	if c.coalescer != nil {
		c.coalescer.Close()
	}
	c.poolMu.Lock()
	for _, b := range c.pool {
		b.Free()
	}
	c.pool = nil
	c.poolMu.Unlock()
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

// Batch is one client's set of page-locked buffers, allocated by C once and reused from call to
// call.  The GPU reads Args/Sel and writes Values across PCIe itself, so what the caller puts in
// the slices is what travels: 16-25 bytes per call in, 8 out - no 48-byte request records, no
// staging copy, and no Go pointer ever reaches C (the slices are views of C memory).
type Batch struct {
	capacity, argsPerCall int
	cArgs, cSel, cValues  unsafe.Pointer // ptype_b200_alloc_pinned
	cReplies              unsafe.Pointer // 16-byte records, only for the array form of CallBatch
	cFailed               unsafe.Pointer // C.malloc: capacity entries, never truncated

	Args   []int64 // [n*argsPerCall]: calculator.Args{A,B} or prime.Args{Min,Max,Target} per call
	Sel    []byte  // [n]: index into the method list of the call (batches naming several methods)
	Values []int64 // [n]: *reply of every call, 0 for a call that failed
	Failed []FailedCall

	retry *Batch // the failed calls of an attempt are re-issued from here
}

// FailedCall mirrors ptype_failed: a call whose status is not StatusOK after the last attempt.
type FailedCall struct {
	Index  uint32
	Status uint32
}

// NewBatch allocates buffers for up to `capacity` calls of argsPerCall (2 or 3) int64 each.
func NewBatch(capacity, argsPerCall int) (*Batch, error) {
	if capacity < 1 || (argsPerCall != 2 && argsPerCall != 3) {
		return nil, errors.New("ptype_b200: bad batch shape")
	}
	b := &Batch{capacity: capacity, argsPerCall: argsPerCall}
	b.cArgs = C.ptype_b200_alloc_pinned(C.size_t(capacity * argsPerCall * 8))
	b.cSel = C.ptype_b200_alloc_pinned(C.size_t(capacity))
	b.cValues = C.ptype_b200_alloc_pinned(C.size_t(capacity * 8))
	b.cFailed = C.malloc(C.size_t(capacity) * C.sizeof_ptype_failed)
	if b.cArgs == nil || b.cSel == nil || b.cValues == nil || b.cFailed == nil {
		b.Free()
		return nil, errors.New("ptype_b200: out of page-locked memory")
	}
	b.Args = (*[1 << 30]int64)(b.cArgs)[: capacity*argsPerCall : capacity*argsPerCall]
	b.Sel = (*[1 << 30]byte)(b.cSel)[:capacity:capacity]
	b.Values = (*[1 << 30]int64)(b.cValues)[:capacity:capacity]
	return b, nil
}

// Free returns the buffers to C.  The slices must not be used afterwards.
func (b *Batch) Free() {
	C.ptype_b200_free_pinned(b.cArgs)
	C.ptype_b200_free_pinned(b.cSel)
	C.ptype_b200_free_pinned(b.cValues)
	C.ptype_b200_free_pinned(b.cReplies)
	C.free(b.cFailed)
	b.cArgs, b.cSel, b.cValues, b.cReplies, b.cFailed = nil, nil, nil, nil, nil
	b.Args, b.Sel, b.Values = nil, nil, nil
	if b.retry != nil {
		b.retry.Free()
		b.retry = nil
	}
}

func (c *BatchClient) serviceKey() C.uint64_t {
	c.gpu.mu.Lock()
	salt := c.gpu.salt
	c.gpu.mu.Unlock()
	cservice := C.CString(c.service)
	defer C.free(unsafe.Pointer(cservice))
	return C.ptype_b200_service_key(cservice, C.uint64_t(salt))
}

func methodRefs(serviceMethods []string) ([C.PTYPE_BATCH_METHODS]C.ptype_method_ref, error) {
	var refs [C.PTYPE_BATCH_METHODS]C.ptype_method_ref
	if len(serviceMethods) < 1 || len(serviceMethods) > C.PTYPE_BATCH_METHODS {
		return refs, errors.New("ptype_b200: a batch names 1 to 4 methods")
	}
	for i, sm := range serviceMethods {
		csm := C.CString(sm)
		C.ptype_b200_method_key(csm, &refs[i].method, &refs[i].recv)
		C.free(unsafe.Pointer(csm))
	}
	return refs, nil
}

// submitValues is one attempt of a values batch: reserve the sequence range with one atomic add
// (rpc.go:181 batched: the first call sees base+1), one C call, collect the failed list.
func (c *BatchClient) submitValues(key C.uint64_t, refs *[C.PTYPE_BATCH_METHODS]C.ptype_method_ref, nm int,
	b *Batch, n int) ([]FailedCall, error) {
	m := uint64(n)
	base := atomic.AddUint64(&c.seq, m) - m
	var sel *C.uint8_t
	if nm > 1 {
		sel = (*C.uint8_t)(b.cSel)
	}
	var count C.uint64_t
	rc := C.ptype_b200_call_batch_values(c.gpu.ctx, 0, key, C.uint64_t(base), &refs[0], C.int(nm), sel,
		(*C.int64_t)(b.cArgs), C.int(b.argsPerCall), C.size_t(n), (*C.int64_t)(b.cValues),
		(*C.ptype_failed)(b.cFailed), C.size_t(b.capacity), &count, 0)
	if err := abiError("ptype_b200_call_batch_values", rc); err != nil {
		return nil, err
	}
	raw := (*[1 << 28]C.ptype_failed)(b.cFailed)[:int(count):int(count)]
	failed := make([]FailedCall, int(count))
	for i, f := range raw {
		failed[i] = FailedCall{Index: uint32(f.index), Status: uint32(f.status)}
	}
	return failed, nil
}

// CallBatchValues issues the first n calls written into b, each naming serviceMethods[b.Sel[j]]
// (b.Sel is ignored when one method is given).  Routing equals n sequential Call()s.  Results
// are b.Values and b.Failed.  Calls that failed are re-issued, under fresh sequence numbers, up
// to cfg.Retries more times (withRetry, rpc.go:107-116, bounded).
func (c *BatchClient) CallBatchValues(serviceMethods []string, b *Batch, n int) error {
	return c.callBatchValues(serviceMethods, b, n, c.cfg.Retries)
}

func (c *BatchClient) callBatchValues(serviceMethods []string, b *Batch, n, retries int) error {
	b.Failed = b.Failed[:0]
	if n == 0 {
		return nil
	}
	if n > b.capacity {
		return errors.New("ptype_b200: batch larger than its buffers")
	}
	refs, err := methodRefs(serviceMethods)
	if err != nil {
		return err
	}
	key := c.serviceKey()
	failed, err := c.submitValues(key, &refs, len(serviceMethods), b, n)
	if err != nil {
		return err
	}
	apc := b.argsPerCall
	for attempt := 0; attempt < retries && len(failed) > 0; attempt++ {
		sort.Slice(failed, func(x, y int) bool { return failed[x].Index < failed[y].Index })
		m := len(failed)
		if b.retry == nil || b.retry.capacity < m {
			if b.retry != nil {
				b.retry.Free()
			}
			if b.retry, err = NewBatch(m, apc); err != nil {
				return err
			}
		}
		r := b.retry
		for t, f := range failed {
			j := int(f.Index)
			copy(r.Args[t*apc:(t+1)*apc], b.Args[j*apc:(j+1)*apc])
			r.Sel[t] = b.Sel[j]
		}
		again, err := c.submitValues(key, &refs, len(serviceMethods), r, m)
		if err != nil {
			return err
		}
		for t, f := range failed {
			b.Values[f.Index] = r.Values[t]
		}
		for i := range again {
			again[i].Index = failed[again[i].Index].Index
		}
		failed = again
	}
	b.Failed = failed
	return nil
}

// CallBatch is the array form: len(args) calls of one method, with the routing decision of every
// call in the reply.  Arguments are copied into a pooled Batch, the GPU writes 16-byte reply
// records, those are unpacked; failed calls go round again like in CallBatchValues.
func (c *BatchClient) CallBatch(serviceMethod string, args [][3]int64, replies []BatchReply) error {
	return c.callBatch(serviceMethod, args, replies, c.cfg.Retries)
}

func (c *BatchClient) callBatch(serviceMethod string, args [][3]int64, replies []BatchReply, retries int) error {
	n := len(args)
	if n == 0 {
		return nil
	}
	if len(replies) < n {
		return errors.New("ptype_b200: replies shorter than args")
	}
	refs, err := methodRefs([]string{serviceMethod})
	if err != nil {
		return err
	}
	key := c.serviceKey()
	apc := 2
	for _, a := range args {
		if a[2] != 0 {
			apc = 3
			break
		}
	}
	b, err := c.acquire(n, apc)
	if err != nil {
		return err
	}
	defer c.release(b)
	if b.cReplies == nil {
		if b.cReplies = C.ptype_b200_alloc_pinned(C.size_t(b.capacity) * C.sizeof_ptype_reply); b.cReplies == nil {
			return errors.New("ptype_b200: out of page-locked memory")
		}
	}
	out := (*[1 << 28]C.ptype_reply)(b.cReplies)[:b.capacity:b.capacity]
	todo := make([]int, n)
	for j := range todo {
		todo[j] = j
	}
	for attempt := 0; ; attempt++ {
		m := uint64(len(todo))
		for t, j := range todo {
			copy(b.Args[t*apc:(t+1)*apc], args[j][:apc])
		}
		base := atomic.AddUint64(&c.seq, m) - m
		rc := C.ptype_b200_call_batch(c.gpu.ctx, 0, key, C.uint64_t(base), &refs[0], 1, nil,
			(*C.int64_t)(b.cArgs), C.int(apc), C.size_t(m), (*C.ptype_reply)(b.cReplies), 0)
		if err := abiError("ptype_b200_call_batch", rc); err != nil {
			return err
		}
		var next []int
		for t, j := range todo {
			replies[j] = BatchReply{Value: int64(out[t].value), Node: uint32(out[t].node), Status: uint32(out[t].status)}
			if out[t].status != StatusOK {
				next = append(next, j)
			}
		}
		if len(next) == 0 || attempt >= retries {
			return nil
		}
		todo = next
	}
}

// Batches behind the array form: Call is safe from many goroutines (rpc.go:59), so every call in
// flight takes its own buffers from a small pool.
func (c *BatchClient) acquire(n, apc int) (*Batch, error) {
	c.poolMu.Lock()
	for i, b := range c.pool {
		if b.capacity >= n && b.argsPerCall == apc {
			c.pool = append(c.pool[:i], c.pool[i+1:]...)
			c.poolMu.Unlock()
			return b, nil
		}
	}
	c.poolMu.Unlock()
	capacity := 64
	for capacity < n {
		capacity *= 2
	}
	return NewBatch(capacity, apc)
}

func (c *BatchClient) release(b *Batch) {
	c.poolMu.Lock()
	if len(c.pool) < 8 {
		c.pool = append(c.pool, b)
		b = nil
	}
	c.poolMu.Unlock()
	if b != nil {
		b.Free()
	}
}

// Go is (*Client).Go (rpc.go:69-105): it returns at once with a *rpc.Call whose Done channel
// receives the call when it has completed.  Like the reference it makes ONE attempt: rpc.go:91
// tests call.Error of the outer call, still nil at that point, so the retry loop never repeats
// (SURVEY.md D6).  A failed call is delivered with Error set and *reply untouched
// (rpc_test.go:262-292).  The call joins the client's Coalescer, so concurrent Go()s and Call()s
// share GPU batches instead of costing a goroutine and a launch each.
func (c *BatchClient) Go(serviceMethod string, args [3]int64, reply *int64, done chan *rpc.Call) *rpc.Call {
	if done == nil {
		done = make(chan *rpc.Call, 10) // rpc.go:72-74
	}
	call := &rpc.Call{ServiceMethod: serviceMethod, Args: args, Reply: reply, Done: done}
	c.coalescerOnce.Do(func() { c.coalescer = NewCoalescer(c) })
	p := &pendingCall{method: serviceMethod, args: args, oneAttempt: true, ready: make(chan struct{})}
	c.coalescer.in <- p
	go func() {
		<-p.ready
		if p.err != nil {
			call.Error = p.err
		} else if err := StatusError(p.rep.Status, serviceMethod); err != nil {
			call.Error = err
		} else {
			*reply = p.rep.Value
		}
		done <- call
	}()
	return call
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
	method     string
	args       [3]int64
	oneAttempt bool // came through Go(): never retried
	rep        BatchReply
	err        error
	ready      chan struct{}
}

type groupKey struct {
	method     string
	oneAttempt bool
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
		groups := map[groupKey][]*pendingCall{{first.method, first.oneAttempt}: {first}}
		n := 1
		timer := time.NewTimer(co.MaxWait)
	gather:
		for n < co.MaxBatch {
			select {
			case p, ok := <-co.in:
				if !ok {
					break gather
				}
				k := groupKey{p.method, p.oneAttempt}
				groups[k] = append(groups[k], p)
				n++
			case <-timer.C:
				break gather
			}
		}
		timer.Stop()
		for k, ps := range groups {
			args := make([][3]int64, len(ps))
			for i, p := range ps {
				args[i] = p.args
			}
			reps := make([]BatchReply, len(ps))
			retries := co.c.cfg.Retries
			if k.oneAttempt {
				retries = 0
			}
			err := co.c.callBatch(k.method, args, reps, retries)
			for i, p := range ps {
				p.rep, p.err = reps[i], err
				close(p.ready)
			}
		}
	}
}
