// Source-only: not compiled in this repository's build image (no Go toolchain). See INTEGRATION.md.
//go:build cgo && cuda

package miner

/*
#cgo LDFLAGS: -lotedama_b200
#include <stdlib.h>
#include <string.h>
#include "otedama_b200.h"
*/
import "C"

import (
	"context"
	"errors"
	"fmt"
	"runtime"
	"sync"
	"sync/atomic"
	"time"
	"unsafe"
)

// Backend is the method set the engine uses on a worker (setup.go:59-77, run.go:239-241,877-879,
// arbitrate.go:255-262, stats.go:123-137). *Worker (CPU) and *GPUWorker both satisfy it; the engine's
// []*miner.Worker becomes []miner.Backend.
type Backend interface {
	Start(ctx context.Context) <-chan Share
	Stop()
	SetWork(*Work)
	HasWork() bool
	DeviceID() string
	Stats() Stats
}

// GPUShare is what a GPU worker knows about a share beyond miner.Share: the extranonce2 value that was
// rolled into the coinbase on the device and the pool's job id string (Stratum V1). StartGPU delivers
// these; Start delivers the embedded miner.Share only, so every existing consumer stays untouched.
type GPUShare struct {
	Share
	ExtraNonce2     uint64
	ExtraNonce2Size int
	JobIDStr        string
	WorkSeq         uint64
}

// GPUWorkExt is the search-space extension of miner.Work (otd_work_t): BIP-320 version rolling and the
// Stratum V1 coinbase with on-device extranonce2 rolling. The zero value means "plain header".
type GPUWorkExt struct {
	VersionMask                              uint32
	VersionBegin, VersionCount, VersionStride uint32
	Coinb1, ExtraNonce1, Coinb2              []byte
	MerkleBranches                           [][32]byte
	ExtraNonce2Size                          uint32
	ExtraNonce2Begin, ExtraNonce2Count, ExtraNonce2Stride uint64
	JobIDStr                                 string
	Repeat                                   bool
}

// GPUWorkerConfig is otd_worker_config_t (config knobs a GPU needs and WorkerConfig has no place for).
type GPUWorkerConfig struct {
	RingCapacity  uint32 // device share ring entries per launch; 0 = 4096
	QueueCapacity uint32 // host share queue; 0 = 65536
	LaunchLog2    uint32 // candidates per kernel launch = 2^LaunchLog2 (16..32); 0 = 32
	KernelVariant int32  // -1 = library default
}

type GPUWorker struct {
	h        *C.otd_worker_t
	deviceID string
	started  atomic.Bool
	cancel   atomic.Pointer[context.CancelFunc]
	done     chan struct{}
	dropped  atomic.Uint64 // shares lost because the Go channel was full (worker.go:271-275 counts the same)
	lastErr  atomic.Pointer[string]
	en2Sizes sync.Map // work_seq (uint64) -> extranonce2_size (uint32) of that work; a share names its work_seq
}

// lastError runs a failing C call and reads otd_last_error() on the SAME OS thread: the message is
// thread-local in the library and a goroutine may migrate between two cgo calls.
func lastError(call func() C.int) (C.int, string) {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	rc := call()
	if rc < 0 {
		return rc, C.GoString(C.otd_last_error())
	}
	return rc, ""
}

// NewGPUWorker fails when the device cannot be opened (no CPU fallback behind it).
func NewGPUWorker(deviceIndex int, deviceID string, cfg GPUWorkerConfig) (*GPUWorker, error) {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	var c C.otd_worker_config_t
	c.ring_capacity, c.queue_capacity = C.uint32_t(cfg.RingCapacity), C.uint32_t(cfg.QueueCapacity)
	c.launch_log2, c.kernel_variant = C.uint32_t(cfg.LaunchLog2), C.int32_t(cfg.KernelVariant)
	h := C.otd_worker_create(C.int(deviceIndex), &c)
	if h == nil {
		return nil, fmt.Errorf("miner: gpu worker %s: %s", deviceID, C.GoString(C.otd_last_error()))
	}
	return &GPUWorker{h: h, deviceID: deviceID, done: make(chan struct{})}, nil
}

// Start mirrors (*Worker).Start (worker.go:127-151); a second call (of Start or StartGPU) panics.
func (w *GPUWorker) Start(ctx context.Context) <-chan Share {
	in := w.StartGPU(ctx)
	out := make(chan Share, 256)
	go func() {
		defer close(out)
		for g := range in {
			select { // non-blocking, counted like worker.go:271-275
			case out <- g.Share:
			default:
				w.dropped.Add(1)
			}
		}
	}()
	return out
}

// StartGPU is Start with the GPU-only share fields: one goroutine parks in otd_worker_poll_shares and
// forwards what the device found.
func (w *GPUWorker) StartGPU(ctx context.Context) <-chan GPUShare {
	if !w.started.CompareAndSwap(false, true) {
		panic("miner: Worker.Start called more than once")
	}
	shares := make(chan GPUShare, 256)
	inner, cancel := context.WithCancel(ctx) // Stop cancels this, like worker.go:131-132
	w.cancel.Store(&cancel)
	go func() {
		defer close(w.done)
		defer close(shares)
		buf := make([]C.otd_share_t, 64)
		backoff := time.Duration(0)
		for inner.Err() == nil {
			rc, msg := lastError(func() C.int {
				return C.otd_worker_poll_shares(w.h, &buf[0], C.int(len(buf)), 100)
			})
			if rc < 0 {
				// a failed run of the current work: reported once by the library; keep the message for
				// Err(), and do not spin while the device stays wedged
				w.lastErr.Store(&msg)
				if backoff < time.Second {
					backoff += 50 * time.Millisecond
				}
				select {
				case <-inner.Done():
				case <-time.After(backoff):
				}
				continue
			}
			backoff = 0
			for i := 0; i < int(rc); i++ {
				s := Share{ChannelID: uint32(buf[i].channel_id), JobID: uint32(buf[i].job_id),
					Nonce: uint32(buf[i].nonce), NTime: uint32(buf[i].ntime), Version: uint32(buf[i].version),
					DeviceID: w.deviceID}
				copy(s.Hash[:], C.GoBytes(unsafe.Pointer(&buf[i].hash[0]), 32))
				size := 0 // of the work the share belongs to, not of whatever work is current by now
				if v, ok := w.en2Sizes.Load(uint64(buf[i].work_seq)); ok {
					size = int(v.(uint32))
				}
				g := GPUShare{Share: s, ExtraNonce2: uint64(buf[i].extranonce2), ExtraNonce2Size: size,
					JobIDStr: C.GoString(&buf[i].job_id_str[0]), WorkSeq: uint64(buf[i].work_seq)}
				select { // non-blocking send, like worker.go:271-275; a full channel is a counted drop
				case shares <- g:
				default:
					w.dropped.Add(1)
				}
			}
		}
	}()
	return shares
}

// Err returns the message of the last failed run (a CUDA error on the worker thread), or nil.
func (w *GPUWorker) Err() error {
	if p := w.lastErr.Load(); p != nil && *p != "" {
		return errors.New(*p)
	}
	return nil
}

// Stop mirrors (*Worker).Stop (worker.go:156-165): cancel, wait for the forwarding
// goroutine to leave otd_worker_poll_shares (at most its 100 ms timeout), then free
// the handle — destroy has to be the last call on it. Safe without Start.
func (w *GPUWorker) Stop() {
	if c := w.cancel.Load(); c != nil {
		(*c)()
		<-w.done
	}
	if w.h != nil {
		C.otd_worker_destroy(w.h)
		w.h = nil
	}
}

// SetWork mirrors (*Worker).SetWork (worker.go:171-176); nil pauses the device.
func (w *GPUWorker) SetWork(work *Work) { w.SetWorkExt(work, nil) }

// SetWorkExt hands the device a work unit with its search-space extension. ext == nil reproduces the
// reference exactly: one header, all 2^32 nonces, rescanned forever.
func (w *GPUWorker) SetWorkExt(work *Work, ext *GPUWorkExt) error {
	if work == nil {
		C.otd_worker_set_work(w.h, nil)
		return nil
	}
	var cw C.otd_work_t // zero value = the reference's nonce-only search of all 2^32 nonces
	cw.job_id, cw.channel_id, cw.nbits = C.uint32_t(work.JobID), C.uint32_t(work.ChannelID), C.uint32_t(work.NBits)
	cw.header.version = C.uint32_t(work.Header.Version)
	cw.header.time, cw.header.bits = C.uint32_t(work.Header.Time), C.uint32_t(work.Header.Bits)
	C.memcpy(unsafe.Pointer(&cw.header.prev_hash[0]), unsafe.Pointer(&work.Header.PrevHash[0]), 32)
	C.memcpy(unsafe.Pointer(&cw.header.merkle_root[0]), unsafe.Pointer(&work.Header.MerkleRoot[0]), 32)
	C.memcpy(unsafe.Pointer(&cw.target[0]), unsafe.Pointer(&work.Target[0]), 32)
	var pin runtime.Pinner // the library copies every buffer before set_work returns
	defer pin.Unpin()
	en2Size := uint32(0)
	if ext == nil {
		cw.flags = C.OTD_WORK_REPEAT // the reference's endless rescan (worker.go:279)
	} else {
		if ext.Repeat {
			cw.flags = C.OTD_WORK_REPEAT
		}
		cw.version_mask = C.uint32_t(ext.VersionMask)
		cw.version_begin, cw.version_count = C.uint32_t(ext.VersionBegin), C.uint32_t(ext.VersionCount)
		cw.version_stride = C.uint32_t(ext.VersionStride)
		if len(ext.JobIDStr) >= C.OTD_JOB_ID_MAX {
			return fmt.Errorf("miner: job id %q longer than %d bytes", ext.JobIDStr, C.OTD_JOB_ID_MAX-1)
		}
		for i := 0; i < len(ext.JobIDStr); i++ {
			cw.job_id_str[i] = C.char(ext.JobIDStr[i])
		}
		if ext.Coinb1 != nil {
			ptr := func(b []byte) (*C.uint8_t, C.uint32_t) {
				if len(b) == 0 {
					return nil, 0
				}
				pin.Pin(&b[0])
				return (*C.uint8_t)(unsafe.Pointer(&b[0])), C.uint32_t(len(b))
			}
			one := []byte{0}
			if len(ext.Coinb1) == 0 {
				pin.Pin(&one[0])
				cw.coinb1 = (*C.uint8_t)(unsafe.Pointer(&one[0])) // non-NULL switches the coinbase path on
			} else {
				cw.coinb1, cw.coinb1_len = ptr(ext.Coinb1)
			}
			cw.extranonce1, cw.extranonce1_len = ptr(ext.ExtraNonce1)
			cw.coinb2, cw.coinb2_len = ptr(ext.Coinb2)
			flat := make([]byte, 0, 32*len(ext.MerkleBranches))
			for i := range ext.MerkleBranches {
				flat = append(flat, ext.MerkleBranches[i][:]...)
			}
			cw.merkle_branches, _ = ptr(flat)
			cw.n_branches = C.uint32_t(len(ext.MerkleBranches))
			cw.extranonce2_size = C.uint32_t(ext.ExtraNonce2Size)
			cw.extranonce2_begin, cw.extranonce2_count = C.uint64_t(ext.ExtraNonce2Begin), C.uint64_t(ext.ExtraNonce2Count)
			cw.extranonce2_stride = C.uint64_t(ext.ExtraNonce2Stride)
		}
		if ext.Coinb1 != nil {
			en2Size = ext.ExtraNonce2Size
		}
	}
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	seq := C.otd_worker_set_work(w.h, &cw)
	if seq < 0 {
		return fmt.Errorf("miner: gpu worker %s: %s", w.deviceID, C.GoString(C.otd_last_error()))
	}
	// shares of this work may still arrive after later SetWork calls; keep a short history
	w.en2Sizes.Store(uint64(seq), en2Size)
	if seq > 64 {
		w.en2Sizes.Delete(uint64(seq) - 64)
	}
	return nil
}

func (w *GPUWorker) HasWork() bool    { return C.otd_worker_has_work(w.h) == 1 }
func (w *GPUWorker) DeviceID() string { return w.deviceID }

func (w *GPUWorker) Stats() Stats { // worker.go:189-212
	var s C.otd_stats_t
	C.otd_worker_stats(w.h, &s)
	return Stats{HashesTotal: uint64(s.hashes_total), SharesFound: uint64(s.shares_found),
		SharesDropped: uint64(s.shares_dropped) + w.dropped.Load(), // the library's drops + the Go channel's
		Uptime:        time.Duration(s.uptime_ns), HashRate: float64(s.hash_rate)}
}
