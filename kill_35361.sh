ssh test@127.0.0.1 "kill -9 5646"
ssh test@127.0.0.1 "kill -9 5647"
kill -9 5044
