ssh test@127.0.0.1 "kill -9 5348"
ssh test@127.0.0.1 "kill -9 5349"
kill -9 5044
