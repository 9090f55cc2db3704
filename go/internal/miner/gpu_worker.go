// Source-only: not compiled in this repository's build image (no Go toolchain). See INTEGRATION.md.
//go:build cgo && cuda

package miner

/*
#cgo LDFLAGS: -lotedama_b200
#include <stdlib.h>
#include "otedama_b200.h"
*/
import "C"

import (
	"context"
	"fmt"
	"sync/atomic"
	"time"
	"unsafe"
)

type GPUWorker struct {
	h        *C.otd_worker_t
	deviceID string
	started  atomic.Bool
	cancel   atomic.Pointer[context.CancelFunc]
	done     chan struct{}
}

// NewGPUWorker fails when the device cannot be opened (no CPU fallback behind it).
func NewGPUWorker(deviceIndex int, deviceID string) (*GPUWorker, error) {
	h := C.otd_worker_create(C.int(deviceIndex), nil)
	if h == nil {
		return nil, fmt.Errorf("miner: gpu worker %s: %s", deviceID, C.GoString(C.otd_last_error()))
	}
	return &GPUWorker{h: h, deviceID: deviceID, done: make(chan struct{})}, nil
}

// Start mirrors (*Worker).Start (worker.go:127-151): one goroutine parks in
// otd_worker_poll_shares and forwards shares; a second call panics.
func (w *GPUWorker) Start(ctx context.Context) <-chan Share {
	if !w.started.CompareAndSwap(false, true) {
		panic("miner: Worker.Start called more than once")
	}
	shares := make(chan Share, 256)
	inner, cancel := context.WithCancel(ctx) // Stop cancels this, like worker.go:131-132
	w.cancel.Store(&cancel)
	go func() {
		defer close(w.done)
		defer close(shares)
		buf := make([]C.otd_share_t, 64)
		for inner.Err() == nil {
			n := int(C.otd_worker_poll_shares(w.h, &buf[0], C.int(len(buf)), 100))
			for i := 0; i < n; i++ {
				s := Share{ChannelID: uint32(buf[i].channel_id), JobID: uint32(buf[i].job_id),
					Nonce: uint32(buf[i].nonce), NTime: uint32(buf[i].ntime), Version: uint32(buf[i].version),
					DeviceID: w.deviceID}
				copy(s.Hash[:], C.GoBytes(unsafe.Pointer(&buf[i].hash[0]), 32))
				select { // non-blocking send, like worker.go:271-275 (the library counts its own drops too)
				case shares <- s:
				default:
				}
			}
		}
	}()
	return shares
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
func (w *GPUWorker) SetWork(work *Work) {
	if work == nil {
		C.otd_worker_set_work(w.h, nil)
		return
	}
	var cw C.otd_work_t // zero value = the reference's nonce-only search of all 2^32 nonces
	cw.job_id, cw.channel_id, cw.nbits = C.uint32_t(work.JobID), C.uint32_t(work.ChannelID), C.uint32_t(work.NBits)
	cw.header.version = C.uint32_t(work.Header.Version)
	cw.header.time, cw.header.bits = C.uint32_t(work.Header.Time), C.uint32_t(work.Header.Bits)
	C.memcpy(unsafe.Pointer(&cw.header.prev_hash[0]), unsafe.Pointer(&work.Header.PrevHash[0]), 32)
	C.memcpy(unsafe.Pointer(&cw.header.merkle_root[0]), unsafe.Pointer(&work.Header.MerkleRoot[0]), 32)
	C.memcpy(unsafe.Pointer(&cw.target[0]), unsafe.Pointer(&work.Target[0]), 32)
	cw.flags = C.OTD_WORK_REPEAT // reproduce the reference's endless rescan; drop it once jobs roll
	C.otd_worker_set_work(w.h, &cw)
}

func (w *GPUWorker) HasWork() bool    { return C.otd_worker_has_work(w.h) == 1 }
func (w *GPUWorker) DeviceID() string { return w.deviceID }

func (w *GPUWorker) Stats() Stats { // worker.go:189-212
	var s C.otd_stats_t
	C.otd_worker_stats(w.h, &s)
	return Stats{HashesTotal: uint64(s.hashes_total), SharesFound: uint64(s.shares_found),
		SharesDropped: uint64(s.shares_dropped), Uptime: time.Duration(s.uptime_ns), HashRate: float64(s.hash_rate)}
}
