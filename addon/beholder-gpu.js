/**
 * beholder-gpu — drop-in for the progress handler of tritonmedia/beholder (index.js:127-155).
 *
 * SOURCE ONLY (Node is absent from the image this repo is tested in); the Python module
 * beholder_b200/handler.py is the same logic, executed by the tests.
 *
 *   const gpu = require('beholder-gpu')({ schema, statusEnum, registry })
 *   amqp.listen('v1.telemetry.progress', gpu.handler(async rmsg => { ...db / trello side effects... }))
 */
'use strict'
const native = require('./build/Release/beholder_napi.node')

const KIND = { string: 1, bytes: 2, int32: 3, enum: 3, uint32: 4, sint32: 5, bool: 6, float: 7, double: 8 }
const ROLE = { key: 1, status: 2, value: 4 }

// Byte offsets inside bh_config (include/beholder_gpu.h), little endian.  tests/test_abi.py compiles the
// header, prints offsetof() for every name below and fails when this table and the C struct disagree.
const OFF = { struct_size: 0, device: 4, mode: 8, ops: 12, n_fields: 36, fields: 40, n_enum: 296, enum_numbers: 300, max_batch_latency_ms: 600, SIZE: 616, FIELD_SIZE: 16 }

function packConfig ({ schema, statusEnum, device = 0, maxBatchLatencyMs = 5 }) {
  const buf = Buffer.alloc(OFF.SIZE)
  buf.writeUInt32LE(OFF.SIZE, OFF.struct_size)
  buf.writeInt32LE(device, OFF.device)
  buf.writeUInt32LE(0, OFF.mode)       // BH_MODE_REFERENCE
  buf.writeUInt32LE(1, OFF.ops)        // COUNT
  const fields = Object.entries(schema)
  buf.writeUInt32LE(fields.length, OFF.n_fields)
  fields.forEach(([name, [number, type]], i) => {
    const o = OFF.fields + i * OFF.FIELD_SIZE
    buf.writeUInt32LE(number, o)
    buf.writeUInt32LE(KIND[type], o + 4)
    buf.writeUInt32LE(name === 'status' ? ROLE.status : (name === 'progress' ? ROLE.value : 0), o + 8)
  })
  const numbers = Object.values(statusEnum)
  buf.writeUInt32LE(numbers.length, OFF.n_enum)
  numbers.forEach((n, i) => buf.writeInt32LE(n, OFF.enum_numbers + i * 4))
  buf.writeUInt32LE(maxBatchLatencyMs, OFF.max_batch_latency_ms)
  return buf
}

module.exports = function beholderGpu (cfg) {
  const ctx = native.create(packConfig(cfg))
  const names = {}
  for (const [name, num] of Object.entries(cfg.statusEnum)) names[num] = name.toLowerCase()
  const pending = []                    // [seq, rmsg] awaiting commit
  let seq = 0n
  let unsubmitted = 0
  const commitEvery = BigInt(cfg.commitEvery || 65536)
  // The broker stops delivering at `prefetch` unacked messages (index.js:43: new AMQP(addr, 100, ...)), so a
  // batch goes to the GPU as soon as half a window is waiting; the library's deadline (maxBatchLatencyMs)
  // covers a queue that goes quiet, and its completion thread reports finished batches through onCommit,
  // which N-API delivers on this thread (napi_threadsafe_function).
  const submitAt = Math.max(1, Math.floor((cfg.prefetch === undefined ? 100 : cfg.prefetch) / 2))

  const ackUpTo = committed => {
    while (pending.length && pending[0][0] <= committed) pending.shift()[1].ack()   // index.js:151,154
  }
  const commitListeners = [ackUpTo]     // the handler's acks, plus one entry per amqpTap
  native.onCommit(ctx, committed => { for (const f of commitListeners) f(committed) })

  const api = {
    handler (sideEffects) {
      return async rmsg => {
        native.push(ctx, rmsg.message.content, Date.now())   // copied before returning
        pending.push([++seq, rmsg])
        if (sideEffects) {
          try { await sideEffects(rmsg) } catch (err) { /* index.js:149-152: warn, still ack */ }
        }
        unsubmitted++
        if (BigInt(pending.length) >= commitEvery) { unsubmitted = 0; ackUpTo(native.flush(ctx)); return }
        if (unsubmitted >= submitAt) { unsubmitted = 0; native.submit(ctx) }
        ackUpTo(native.poll(ctx))
      }
    },
    commit () { unsubmitted = 0; ackUpTo(native.flush(ctx)) },
    // Raw-socket ingest (SURVEY section 8 f1): instead of amqp.listen + one rmsg per delivery, hand the tap
    // every chunk the broker socket emits.  Bodies go to the GPU ring in one batch per chunk; `ack` is the
    // batched form of index.js:154 -- one basic.ack with allUpTo for everything the GPU has committed.
    //   const tap = gpu.amqpTap({ consumerTag, channel })        // channel: an amqplib Channel
    //   socket.on('data', chunk => tap.onData(chunk))
    // Delivery tags count per CHANNEL, so the tap keeps only deliveries that arrived on `channel` (channel.ch,
    // amqplib's channel number) with `consumerTag`, and acks on that channel alone.  dedicated = true (default):
    // the progress consumer is the only consumer on its channel, and one basic.ack with allUpTo covers a whole
    // committed batch.  dedicated = false: other consumers share the channel (the reference's status listener,
    // index.js:62, whose messages must stay unacked when its handler throws), so every tag is acked by itself.
    amqpTap ({ consumerTag, channel, dedicated = true }) {
      const tap = native.amqpCreate(consumerTag || '')
      const channelNumber = channel.ch
      let rest = Buffer.alloc(0)
      const waiting = []                 // [seq after this delivery, deliveryTag], in arrival order
      const toNumber = tag => {          // amqplib carries delivery tags as JS numbers
        if (tag > BigInt(Number.MAX_SAFE_INTEGER)) throw new RangeError('delivery tag beyond 2^53: reopen the channel')
        return Number(tag)
      }
      const ackCommittedUpTo = committed => {
        let last = null
        while (waiting.length && waiting[0][0] <= committed) {
          last = waiting.shift()[1]
          if (!dedicated) channel.ack({ fields: { deliveryTag: toNumber(last) } }, false)
        }
        if (dedicated && last !== null) channel.ack({ fields: { deliveryTag: toNumber(last) } }, true)
      }
      commitListeners.push(ackCommittedUpTo)
      return {
        onData (chunk) {
          const buf = rest.length ? Buffer.concat([rest, chunk]) : chunk
          // bodies of deliveries on other channels are not pushed (the shim filters on channelNumber)
          const { consumed, tags } = native.amqpFeed(tap, ctx, buf, Date.now(), channelNumber)
          rest = buf.subarray(consumed)
          for (const tag of tags) waiting.push([++seq, tag])
          ackCommittedUpTo(native.poll(ctx))
        },
        // acks also arrive through the completion callback; by hand e.g. before closing the channel
        ackCommitted (flush = false) { ackCommittedUpTo(flush ? native.flush(ctx) : native.poll(ctx)) }
      }
    },
    // { '<lower-cased enum name>': count } — the values of beholder_progress_updates_total{status=...}
    statusCounts () {
      api.commit()
      const { rows, data } = native.snapshot(ctx, 0n, 64)
      const out = {}
      for (let i = 0; i < Number(rows); i++) {
        const status = Number(data.readBigUInt64LE(i * 48))
        if (names[status] !== undefined) out[names[status]] = Number(data.readBigUInt64LE(i * 48 + 16))
      }
      return out
    }
  }

  // prom-client custom collector under the reference's metric name (index.js:30-34)
  if (cfg.registry) {
    cfg.registry.registerMetric({
      name: 'beholder_progress_updates_total',
      get () {
        return {
          name: 'beholder_progress_updates_total',
          help: 'Total number of messages processed in this processes lifetime',
          type: 'counter',
          aggregator: 'sum',
          values: Object.entries(api.statusCounts()).map(([status, value]) => ({ value, labels: { status } }))
        }
      }
    })
  }
  return api
}
