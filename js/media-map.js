// Same module path and default export as lib/integration/mapping/media-map.js
module.exports = require('./mapping').MediaMap;
module.exports.default = module.exports;
