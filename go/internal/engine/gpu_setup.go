// Source-only: not compiled in this repository's build image (no Go toolchain). See INTEGRATION.md.
//go:build cgo && cuda

package engine

import (
	"context"
	"fmt"

	"github.com/shizukutanaka/Otedama/internal/config"
	"github.com/shizukutanaka/Otedama/internal/hal"
	"github.com/shizukutanaka/Otedama/internal/miner"
)

// startGPUWorkers is startMinerWorkers (setup.go:56-78) for devices of the CUDA driver: one GPUWorker per
// SHA256d-capable CUDA device that the allow-list lets through, configured per device, plus the merged
// share channel. CPU devices keep going through startMinerWorkers; the engine appends both worker sets
// into one []miner.Backend and fans both share channels in with mergeShares (fanin.go:66).
//
// Differences from the reference, all of them additions:
//   - gpu.devices allow-list: a device that is not listed gets no worker (and a log line saying so); a
//     listed ID that was not detected is an error;
//   - gpu.defaults / gpu.per_device: ring, host queue, launch size and kernel variant reach
//     otd_worker_create for that device only.
func startGPUWorkers(ctx context.Context, devices []hal.Device, gpu config.GPUConfig,
	log func(level, msg string)) ([]*miner.GPUWorker, <-chan miner.GPUShare, error) {
	detected := map[string]bool{}
	for _, dev := range devices {
		detected[dev.Identity().ID] = true
	}
	for _, id := range gpu.Devices {
		if !detected[id] {
			return nil, nil, fmt.Errorf("engine: gpu.devices names %q, which was not detected", id)
		}
	}
	for id := range gpu.PerDevice {
		if !detected[id] {
			return nil, nil, fmt.Errorf("engine: gpu.per_device has settings for %q, which was not detected", id)
		}
	}
	var workers []*miner.GPUWorker
	var chans []<-chan miner.GPUShare
	stopAll := func() {
		for _, w := range workers {
			w.Stop()
		}
	}
	for _, dev := range devices {
		idx, isCUDA := dev.(hal.CUDAIndexed)
		if !isCUDA || !dev.Capabilities().SHA256d {
			continue
		}
		id := dev.Identity().ID
		if !gpu.Allowed(id) {
			log("info", fmt.Sprintf("engine: %s not in gpu.devices, no worker", id))
			continue
		}
		dc := gpu.For(id)
		variant := int32(-1)
		if dc.KernelVariant != nil {
			variant = *dc.KernelVariant
		}
		w, err := miner.NewGPUWorker(idx.Index(), id, miner.GPUWorkerConfig{RingCapacity: dc.RingCapacity,
			QueueCapacity: dc.QueueCapacity, LaunchLog2: dc.LaunchLog2, KernelVariant: variant})
		if err != nil {
			stopAll()
			return nil, nil, err
		}
		workers = append(workers, w)
		chans = append(chans, w.StartGPU(ctx))
		log("info", fmt.Sprintf("engine: gpu worker for %s", dev.Identity()))
	}
	if len(workers) == 0 {
		return nil, nil, fmt.Errorf("engine: no SHA256d-capable CUDA devices found")
	}
	return workers, fanIn(ctx, chans, 4), nil // the generic fan-in behind mergeShares (fanin.go:22-63)
}

// pauseGPUDevice is applyAllocation's pauseDevice (arbitrate.go:255-262) over GPU workers: SetWork(nil) on
// the one worker whose DeviceID matches; every other device keeps hashing.
func pauseGPUDevice(workers []*miner.GPUWorker, deviceID string) bool {
	for _, w := range workers {
		if w.DeviceID() == deviceID {
			w.SetWork(nil)
			return true
		}
	}
	return false
}

// setPartitionedWork hands each of the G GPU workers its slice of one job's search space — the static,
// zero-communication partition of SURVEY §8e (otd_work_partition does the same for C callers):
//   - BIP-320 rolled work: version index k -> device k mod G  (VersionBegin = g, VersionStride = G);
//   - Stratum V1 coinbase work: extranonce2 value e -> device e mod G;
//   - a plain header: G contiguous nonce windows are not expressible through SetWorkExt (the reference's
//     Work has no nonce window) — a plain header goes to every device only when G == 1.
func setPartitionedWork(workers []*miner.GPUWorker, work *miner.Work, ext miner.GPUWorkExt) error {
	G := uint32(len(workers))
	for g, w := range workers {
		e := ext
		switch {
		case ext.VersionMask != 0:
			stride := ext.VersionStride
			if stride == 0 {
				stride = 1
			}
			e.VersionBegin = ext.VersionBegin + uint32(g)*stride
			e.VersionStride = stride * G
			if ext.VersionCount != 0 {
				e.VersionCount = (ext.VersionCount - uint32(g) + G - 1) / G
			}
		case ext.Coinb1 != nil:
			stride := ext.ExtraNonce2Stride
			if stride == 0 {
				stride = 1
			}
			e.ExtraNonce2Begin = ext.ExtraNonce2Begin + uint64(g)*stride
			e.ExtraNonce2Stride = stride * uint64(G)
			e.ExtraNonce2Count = (ext.ExtraNonce2Count - uint64(g) + uint64(G) - 1) / uint64(G)
		default:
			if G > 1 {
				return fmt.Errorf("engine: a plain header cannot be split over %d GPUs; roll versions or extranonce2", G)
			}
		}
		if err := w.SetWorkExt(work, &e); err != nil {
			return err
		}
	}
	return nil
}
