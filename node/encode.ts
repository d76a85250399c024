// Drop-in for brotli-lib's src/encode/encode.ts over the B200 hot path: same exports (brotliEncode, BrotliEncoder,
// EncoderMode, BrotliEncodeOptions), same option handling; the work is done by libbrotligpu.so through the Node-API
// addon (node/addon.cc).  There is no CPU fallback: without the addon or a CUDA device the calls throw.
//
// Reference lines mirrored here: option defaults and clamping encode.ts:57-71 (never reject), EncoderMode
// enc-constants.ts:56-60 (TEXT == GENERIC; FONT acts at quality >= 4), fresh caller-owned Uint8Array results
// (bit-writer.ts:146-150), update() returning the bytes completed so far, possibly none (encode.ts:354-388).
// The host-only cases (empty input -> A1 01, quality 0 or < 64 bytes -> uncompressed metablocks, encode.ts:92-138)
// are produced inside the C ABI on the host, byte-identical to the reference.
import { createRequire } from 'node:module'

interface Addon {
  encode(input: Uint8Array, quality: number, lgwin: number, mode: number): Uint8Array
  encodeBatch(inputs: Uint8Array[], quality: number, lgwin: number, mode: number): Uint8Array[]
  encodePart(part: Uint8Array, prev2: Uint8Array, first: boolean, last: boolean, quality: number, lgwin: number, mode: number): Uint8Array
  init(deviceIds: number[]): void
  Encoder: new (quality: number, lgwin: number, mode: number) => { update(chunk: Uint8Array): Uint8Array; finish(): Uint8Array }
}

const addon: Addon = createRequire(import.meta.url)('./build/Release/brotligpu.node')

export const enum EncoderMode {
  GENERIC = 0,
  TEXT = 1,
  FONT = 2,
}

export interface BrotliEncodeOptions {
  quality?: number
  lgwin?: number
  mode?: EncoderMode
  sizeHint?: number   // accepted and ignored, as in the reference (only dead code reads it)
}

function clampOptions(options: BrotliEncodeOptions): [number, number, number] {
  const quality = Math.max(0, Math.min(11, options.quality ?? 11))   // encode.ts:57-62
  const lgwin = Math.max(10, Math.min(24, options.lgwin ?? 22))
  return [quality | 0, lgwin | 0, (options.mode ?? EncoderMode.GENERIC) | 0]
}

export function brotliEncode(input: Uint8Array, options: BrotliEncodeOptions = {}): Uint8Array {
  const [quality, lgwin, mode] = clampOptions(options)
  return addon.encode(input, quality, lgwin, mode)
}

/** Independent inputs as one device job (BASELINE config 5); item i equals brotliEncode(inputs[i], options). */
export function brotliEncodeBatch(inputs: Uint8Array[], options: BrotliEncodeOptions = {}): Uint8Array[] {
  const [quality, lgwin, mode] = clampOptions(options)
  return addon.encodeBatch(inputs, quality, lgwin, mode)
}

/** Use several GPUs of the box for every later brotliEncode / brotliEncodeBatch call (one stitched stream per input). */
export function useDevices(deviceIds: number[]): void {
  addon.init(deviceIds)
}

export class BrotliEncoder {   // encode.ts:290-409
  #h: InstanceType<Addon['Encoder']>
  constructor(options: BrotliEncodeOptions = {}) {
    const [quality, lgwin, mode] = clampOptions(options)
    this.#h = new addon.Encoder(quality, lgwin, mode)
  }
  update(chunk: Uint8Array): Uint8Array {
    return this.#h.update(chunk)
  }
  finish(): Uint8Array {
    return this.#h.finish()
  }
}
