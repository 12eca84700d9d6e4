ssh test@127.0.0.1 "kill -9 5452"
ssh test@127.0.0.1 "kill -9 5453"
ssh test@127.0.0.1 "kill -9 5454"
kill -9 5044
