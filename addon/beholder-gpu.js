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

// packs a bh_config (include/beholder_gpu.h); offsets follow the C struct layout (little endian)
function packConfig ({ schema, statusEnum, device = 0 }) {
  const buf = Buffer.alloc(600)
  buf.writeUInt32LE(600, 0)            // struct_size
  buf.writeInt32LE(device, 4)
  buf.writeUInt32LE(0, 8)              // mode = BH_MODE_REFERENCE
  buf.writeUInt32LE(1, 12)             // ops  = COUNT
  const fields = Object.entries(schema)
  buf.writeUInt32LE(fields.length, 36) // n_fields
  fields.forEach(([name, [number, type]], i) => {
    const o = 40 + i * 16
    buf.writeUInt32LE(number, o)
    buf.writeUInt32LE(KIND[type], o + 4)
    buf.writeUInt32LE(name === 'status' ? ROLE.status : (name === 'progress' ? ROLE.value : 0), o + 8)
  })
  const numbers = Object.values(statusEnum)
  buf.writeUInt32LE(numbers.length, 296) // n_enum
  numbers.forEach((n, i) => buf.writeInt32LE(n, 300 + i * 4))
  return buf
}

module.exports = function beholderGpu (cfg) {
  const ctx = native.create(packConfig(cfg))
  const names = {}
  for (const [name, num] of Object.entries(cfg.statusEnum)) names[num] = name.toLowerCase()
  const pending = []                    // [seq, rmsg] awaiting commit
  let seq = 0n
  const commitEvery = BigInt(cfg.commitEvery || 65536)

  const ackUpTo = committed => {
    while (pending.length && pending[0][0] <= committed) pending.shift()[1].ack()   // index.js:151,154
  }

  const api = {
    handler (sideEffects) {
      return async rmsg => {
        native.push(ctx, rmsg.message.content, Date.now())   // copied before returning
        pending.push([++seq, rmsg])
        if (sideEffects) {
          try { await sideEffects(rmsg) } catch (err) { /* index.js:149-152: warn, still ack */ }
        }
        ackUpTo(BigInt(pending.length) >= commitEvery ? native.flush(ctx) : native.poll(ctx))
      }
    },
    commit () { ackUpTo(native.flush(ctx)) },
    // Raw-socket ingest (SURVEY section 8 f1): instead of amqp.listen + one rmsg per delivery, hand the tap
    // every chunk the broker socket emits.  Bodies go to the GPU ring in one batch per chunk; `ack` is the
    // batched form of index.js:154 -- one basic.ack with allUpTo for everything the GPU has committed.
    //   const tap = gpu.amqpTap({ consumerTag, channel })        // channel: an amqplib Channel
    //   socket.on('data', chunk => tap.onData(chunk))
    amqpTap ({ consumerTag, channel }) {
      const tap = native.amqpCreate(consumerTag || '')
      let rest = Buffer.alloc(0)
      let lastTag = 0n
      let acked = 0n
      return {
        onData (chunk) {
          const buf = rest.length ? Buffer.concat([rest, chunk]) : chunk
          const { consumed, tags } = native.amqpFeed(tap, ctx, buf, Date.now())
          rest = buf.subarray(consumed)
          if (tags.length) lastTag = tags[tags.length - 1]
        },
        // call on a timer or every N chunks: everything pushed so far is aggregated after flush()
        ackCommitted () {
          native.flush(ctx)
          if (lastTag > acked) { channel.ack({ fields: { deliveryTag: Number(lastTag) } }, true); acked = lastTag }
        }
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
